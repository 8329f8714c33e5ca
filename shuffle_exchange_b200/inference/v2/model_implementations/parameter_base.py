"""Declarative checkpoint parameters (reference: ``inference/v2/model_implementations/parameter_base.py``).

A :class:`ParameterBase` subclass names, as class annotations, the raw checkpoint tensors it is assembled from::

    class UnfusedQKVParameter(ParameterBase):
        q_params: torch.Tensor
        k_params: torch.Tensor
        v_params: torch.Tensor
        def finalize(self): return torch.cat([self.q_params, self.k_params, self.v_params], dim=0)

``ParamList("n_experts")`` declares a list dependency whose length is read from the model (``model.n_experts``).  The
owning :class:`LayerContainer` feeds dependencies in whatever order the checkpoint yields them; as soon as the last
one arrives ``finalize()`` runs and the result is stored on the container.  (The reference generates properties with
a metaclass; here ``__init_subclass__`` collects the annotations and ``__setattr__`` does the bookkeeping.)"""
from __future__ import annotations

from typing import Any, Dict, List, Optional

import torch


class _ListSpec:
    def __init__(self, count_attr: str):
        self.count_attr = count_attr


def ParamList(count_attr: str) -> _ListSpec:
    """Annotation for a dependency that is a list of tensors, ``getattr(model, count_attr)`` long."""
    return _ListSpec(count_attr)


class ParametrizedList:
    """Fixed-length list of tensors that reports when every slot has been filled."""

    def __init__(self, owner: "ParameterBase", n_params: int):
        self._owner, self.n_params = owner, int(n_params)
        self._items: List[Optional[torch.Tensor]] = [None] * self.n_params
        self.set_params = 0

    def __getitem__(self, i):
        return self._items[i]

    def __setitem__(self, i, value):
        if self._items[i] is not None:
            raise ValueError(f"element {i} of a parameter list was set twice")
        self._items[i] = value
        self.set_params += 1
        self._owner._dependency_arrived()

    def __iter__(self):
        return iter(self._items)

    def __len__(self):
        return self.n_params

    @property
    def complete(self) -> bool:
        return self.set_params == self.n_params


class ParameterBase:
    _tensor_deps: tuple = ()
    _list_deps: Dict[str, str] = {}

    def __init_subclass__(cls, **kw):
        super().__init_subclass__(**kw)
        tensors, lists = list(cls._tensor_deps), dict(cls._list_deps)
        for name, ann in cls.__dict__.get("__annotations__", {}).items():
            if isinstance(ann, _ListSpec):
                lists[name] = ann.count_attr
            elif ann is torch.Tensor or ann == "torch.Tensor":
                tensors.append(name)
        cls._tensor_deps, cls._list_deps = tuple(dict.fromkeys(tensors)), lists

    def __init__(self, model: Any, parent_container: Any = None) -> None:
        object.__setattr__(self, "_ready", False)
        self.inference_model = model
        self.completed_components = 0
        self._parent = parent_container
        self._name_in_parent: Optional[str] = None
        self._finalized = False
        for n in self._tensor_deps:
            object.__setattr__(self, n, None)
        for n, count_attr in self._list_deps.items():
            object.__setattr__(self, n, ParametrizedList(self, getattr(model, count_attr)))
        object.__setattr__(self, "_ready", True)

    def __setattr__(self, name, value):
        if getattr(self, "_ready", False) and name in self._tensor_deps:
            if getattr(self, name) is not None:
                raise ValueError(f"dependency {name} of {type(self).__name__} was set twice")
            object.__setattr__(self, name, value)
            self._dependency_arrived()
            return
        if getattr(self, "_ready", False) and name in self._list_deps:
            raise ValueError(f"{name} is a parameter list: assign its elements")
        object.__setattr__(self, name, value)

    # ---------------------------------------------------------------------------------------------
    @property
    def n_dependencies(self) -> int:
        return len(self._tensor_deps) + sum(getattr(self, n).n_params for n in self._list_deps)

    @property
    def is_complete(self) -> bool:
        return all(getattr(self, n) is not None for n in self._tensor_deps) and \
            all(getattr(self, n).complete for n in self._list_deps)

    def _dependency_arrived(self) -> None:
        self.completed_components += 1
        if self.is_complete and not self._finalized:
            self._finalized = True
            self.complete_component()

    def complete_component(self) -> None:
        value = self.finalize()
        if self._parent is not None and self._name_in_parent is not None:
            self._parent._parameter_finalized(self._name_in_parent, value)
        else:
            self.value = value

    def finalize(self) -> torch.Tensor:
        """Assemble the final tensor from the dependencies."""
        raise NotImplementedError
