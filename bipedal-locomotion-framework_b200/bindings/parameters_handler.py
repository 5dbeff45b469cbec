"""bindings/python/ParametersHandler/src/Module.cpp: ``StdParametersHandler`` with typed setters / getters and groups."""
from __future__ import annotations

import copy
from datetime import timedelta

import numpy as np


class IParametersHandler:
    pass


class StdParametersHandler(IParametersHandler):
    def __init__(self):
        self._values = {}
        self._groups = {}

    # ---- setters (Module.cpp: set_parameter_<type>)
    def set_parameter_bool(self, name, value):
        self._values[name] = bool(value)

    def set_parameter_int(self, name, value):
        self._values[name] = int(value)

    def set_parameter_float(self, name, value):
        self._values[name] = float(value)

    def set_parameter_string(self, name, value):
        self._values[name] = str(value)

    def set_parameter_datetime(self, name, value):
        if not isinstance(value, timedelta):
            raise TypeError("set_parameter_datetime expects a datetime.timedelta")
        self._values[name] = value

    def set_parameter_vector_bool(self, name, value):
        self._values[name] = [bool(v) for v in value]

    def set_parameter_vector_int(self, name, value):
        self._values[name] = [int(v) for v in value]

    def set_parameter_vector_float(self, name, value):
        self._values[name] = np.asarray(value, dtype=np.float64).reshape(-1).copy()

    def set_parameter_vector_string(self, name, value):
        self._values[name] = [str(v) for v in value]

    # ---- getters: raise ValueError when the parameter is missing or of another type, as the reference bindings do
    def _get(self, name, kind, what):
        v = self._values.get(name)
        if not isinstance(v, kind) or (kind is int and isinstance(v, bool)):
            raise ValueError(f"Failed to find a parameter '{name}' of type {what}")
        return v

    def get_parameter_bool(self, name):
        return self._get(name, bool, "bool")

    def get_parameter_int(self, name):
        return self._get(name, int, "int")

    def get_parameter_float(self, name):
        v = self._values.get(name)
        if isinstance(v, bool) or not isinstance(v, (int, float)):
            raise ValueError(f"Failed to find a parameter '{name}' of type float")
        return float(v)

    def get_parameter_string(self, name):
        return self._get(name, str, "string")

    def get_parameter_datetime(self, name):
        v = self._values.get(name)
        if isinstance(v, timedelta):
            return v
        if isinstance(v, float):   # seconds, as the C++ handler converts a double (Core.h)
            return timedelta(seconds=v)
        raise ValueError(f"Failed to find a parameter '{name}' of type datetime")

    def get_parameter_vector_bool(self, name):
        v = self._get(name, list, "vector<bool>")
        if not all(isinstance(x, bool) for x in v):
            raise ValueError(f"Failed to find a parameter '{name}' of type vector<bool>")
        return v

    def get_parameter_vector_int(self, name):
        return self._get(name, list, "vector<int>")

    def get_parameter_vector_float(self, name):
        v = self._values.get(name)
        if isinstance(v, np.ndarray):
            return v
        if isinstance(v, list) and all(isinstance(x, (int, float)) and not isinstance(x, bool) for x in v):
            return np.asarray(v, dtype=np.float64)
        raise ValueError(f"Failed to find a parameter '{name}' of type vector<float>")

    def get_parameter_vector_string(self, name):
        v = self._get(name, list, "vector<string>")
        if not all(isinstance(x, str) for x in v):
            raise ValueError(f"Failed to find a parameter '{name}' of type vector<string>")
        return v

    # ---- groups
    def set_group(self, name, new_group):
        self._groups[name] = copy.deepcopy(new_group)   # StdImplementation::setGroup clones
        return True

    def get_group(self, name):
        if name not in self._groups:
            raise ValueError(f"Failed to find the group '{name}'")
        return self._groups[name]

    def has(self, name):
        return name in self._values

    def is_empty(self):
        return not self._values and not self._groups

    def clear(self):
        self._values.clear()
        self._groups.clear()

    def __repr__(self):
        return "StdParametersHandler(" + ", ".join(list(self._values) + [f"[{g}]" for g in self._groups]) + ")"

    # ---- helpers of this package (no counterpart): typed lookups that return None instead of raising
    def _opt(self, getter, name):
        try:
            return getter(name)
        except ValueError:
            return None
