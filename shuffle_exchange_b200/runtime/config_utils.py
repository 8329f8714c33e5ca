"""Config helpers: pydantic base model with deprecated-field migration, scalar getters, duplicate-key
detection (reference: deepspeed/runtime/config_utils.py:17-118, :205)."""
from __future__ import annotations

import collections
import json
from functools import reduce
from typing import Any, Dict

from pydantic import BaseModel, ConfigDict

from ..utils.logging import logger


class DeepSpeedConfigModel(BaseModel):
    """Base of every structured config block.

    A field may declare ``json_schema_extra={"deprecated": True, "new_param": "a.b", "new_param_fn": f,
    "set_new_param": True}``; when the old key is present its value is forwarded (through
    ``new_param_fn``) to the replacement unless that one is set explicitly, and a warning is logged.
    """

    model_config = ConfigDict(validate_default=True, validate_assignment=True, use_enum_values=True,
                              populate_by_name=True, extra="forbid", arbitrary_types_allowed=True,
                              protected_namespaces=())

    def __init__(self, strict: bool = False, **data):
        if not strict:  # drop "auto" placeholders so defaults apply (HF integration convention)
            data = {k: v for k, v in data.items() if not (v == "auto" and k != "replace_method")}
        super().__init__(**data)
        self._deprecated_fields_check()

    def _process_deprecated_field(self, dep_field: str):
        fields_set = self.model_fields_set
        info = type(self).model_fields[dep_field]
        kwargs = info.json_schema_extra or {}
        new_param_fn = kwargs.get("new_param_fn", lambda x: x)
        param_value = new_param_fn(getattr(self, dep_field))
        new_field = kwargs.get("new_param", "")
        dep_msg = kwargs.get("deprecated_msg", "")
        if dep_field in fields_set:
            logger.warning(f"Config parameter {dep_field} is deprecated" +
                           (f" use {new_field} instead" if new_field else "") + (f". {dep_msg}" if dep_msg else ""))
            if new_field and kwargs.get("set_new_param", True):
                base = new_field.split(".")[0]
                assert base not in fields_set, f"Cannot provide deprecated parameter '{dep_field}' and replacing parameter '{new_field}' together"
                try:
                    parts = new_field.split(".")
                    obj = reduce(getattr, parts[:-1], self)
                    setattr(obj, parts[-1], param_value)
                except Exception as e:
                    logger.error(f"Tried setting value for '{new_field}' with value from deprecated '{dep_field}'")
                    raise e

    def _deprecated_fields_check(self):
        for name, info in type(self).model_fields.items():
            if (info.json_schema_extra or {}).get("deprecated", False):
                self._process_deprecated_field(name)


def get_config_default(config, field_name):
    assert field_name in config.model_fields, f"'{field_name}' is not a field in {config}"
    assert not config.model_fields.get(field_name).is_required(), f"'{field_name}' is a required field and does not have a default value"
    return config.model_fields.get(field_name).get_default()


class pp_int(int):
    """Int that prints with thousands separators in config dumps (e.g. ``5e8`` -> ``500,000,000``)."""

    def __new__(cls, val, custom_print_str=None):
        inst = super().__new__(cls, val)
        inst.custom_print_str = custom_print_str
        return inst

    def __repr__(self):
        return self.custom_print_str or f"{self.real:,}"


class ScientificNotationEncoder(json.JSONEncoder):
    """JSON encoder used for config dumps: numbers above 1000 are printed in scientific notation
    (``5.000000e+08``) so bucket sizes stay readable."""

    def _fmt(self, obj, depth):
        pad = " " * ((self.indent or 4) * (depth + 1))
        pad_close = " " * ((self.indent or 4) * depth)
        if isinstance(obj, bool):
            return "true" if obj else "false"
        if isinstance(obj, (int, float)):
            return f"{obj:e}" if obj > 1e3 else str(obj)
        if isinstance(obj, dict):
            body = ", ".join(f'\n{pad}"{k}": {self._fmt(v, depth + 1)}' for k, v in obj.items())
            return "{" + body + f"\n{pad_close}" + "}"
        if isinstance(obj, (list, tuple)):
            return "[" + ", ".join(self._fmt(v, depth) for v in obj) + "]"
        return json.dumps(obj, default=str)

    def iterencode(self, o, _one_shot=False):
        yield self._fmt(o, 0)

    def encode(self, o):
        return self._fmt(o, 0)


def get_scalar_param(param_dict: Dict[str, Any], param_name: str, param_default_value):
    return param_dict.get(param_name, param_default_value)


def get_list_param(param_dict, param_name, param_default_value):
    return param_dict.get(param_name, param_default_value)


def get_dict_param(param_dict, param_name, param_default_value):
    return param_dict.get(param_name, param_default_value)


def dict_raise_error_on_duplicate_keys(ordered_pairs):
    """``object_pairs_hook`` rejecting JSON objects that repeat a key."""
    d = dict((k, v) for k, v in ordered_pairs)
    if len(d) != len(ordered_pairs):
        counter = collections.Counter([pair[0] for pair in ordered_pairs])
        keys = [key for key, value in counter.items() if value > 1]
        raise ValueError(f"Duplicate keys in DeepSpeed config: {keys}")
    return d


class DeepSpeedConfigObject:
    """Plain attribute bag with a JSON repr — base of the older, hand-parsed config sections (reference
    ``config_utils.py:176-187``)."""

    def repr(self):
        return self.__dict__

    def __repr__(self):
        return json.dumps(self.__dict__, sort_keys=True, indent=4, cls=ScientificNotationEncoder)
