"""bindings/python/System/src/VariablesHandler.cpp, ConstantWeightProvider.cpp."""
from __future__ import annotations

import numpy as np

from ._log import log


class VariableDescription:
    def __init__(self, offset=-1, size=-1, name=""):
        self.offset, self.size, self.name = offset, size, name

    def is_valid(self):
        return self.offset >= 0 and self.size >= 0

    def __repr__(self):
        return f"VariableDescription(name={self.name!r}, offset={self.offset}, size={self.size})"


class VariablesHandler:
    """src/System/src/VariablesHandler.cpp: named slices of the optimisation vector, in insertion order."""

    def __init__(self):
        self._vars = {}
        self._n = 0

    def add_variable(self, name, size, elements_name=None):
        if name in self._vars:
            log().error("[VariablesHandler::addVariable] The variable named %s already exists.", name)
            return False
        if elements_name is not None and len(elements_name) != size:
            log().error("[VariablesHandler::addVariable] The size of the vector of the element is different from the expected one.")
            return False
        self._vars[name] = VariableDescription(self._n, int(size), name)
        self._n += int(size)
        return True

    def get_variable(self, name):
        return self._vars.get(name, VariableDescription())

    def get_number_of_variables(self):
        return self._n

    def initialize(self, handler):
        """keys: variables_name, variables_size (VariablesHandler.cpp:74-111)"""
        try:
            names, sizes = handler.get_parameter_vector_string("variables_name"), handler.get_parameter_vector_int("variables_size")
        except ValueError as e:
            log().error("[VariablesHandler::initialize] %s", e)
            return False
        if len(names) != len(sizes):
            log().error("[VariablesHandler::initialize] The size of the list of the variables name is different from the size of the variables size.")
            return False
        self._vars, self._n = {}, 0
        return all(self.add_variable(n, s) for n, s in zip(names, sizes))

    def __str__(self):
        return " ".join(f"{v.name} size: {v.size}, offset: {v.offset}." for v in self._vars.values())


class WeightProvider:
    def get_output(self):
        raise NotImplementedError

    def is_output_valid(self):
        return False


class ConstantWeightProvider(WeightProvider):
    def __init__(self, weight=None):
        self.weight = None if weight is None else np.asarray(weight, dtype=np.float64).reshape(-1).copy()

    def initialize(self, handler):
        try:
            self.weight = np.asarray(handler.get_parameter_vector_float("weight"), dtype=np.float64).copy()
        except ValueError:
            log().error("[ConstantWeightProvider::initialize] Unable to get the parameter named 'weight'.")
            return False
        return True

    def get_output(self):
        return self.weight

    def is_output_valid(self):
        return self.weight is not None and self.weight.size > 0
