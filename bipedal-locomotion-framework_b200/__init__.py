"""B200-native batched whole-body inverse kinematics (the IntegrationBasedIK tick of
bipedal-locomotion-framework) -- Python host side.

Import as ``blf_b200`` (the directory name is not a valid Python identifier; the
top-level ``blf_b200.py`` shim maps the name onto this directory).
"""
from . import models, tasks, workloads  # noqa: F401

__all__ = ["models", "tasks", "workloads"]
