"""Operator-provider registry with the reference's contract (operators/registry.py:134-194):

a provider is a class with ``name``, ``priority`` and ``supports(spec, caps) -> SupportResult``;
``OpResolver.resolve`` instantiates the supported provider of highest priority (ties by name),
raises when none supports the spec, and records the binding so that
``operator_runtime_stats()`` (registry.py:31-72) can aggregate ``runtime_kernel_stats()``.
``DeviceCaps`` mirrors platforms/interface.py:27-39.
"""
from __future__ import annotations

import enum
import weakref
from dataclasses import dataclass
from typing import Generic, TypeVar

import torch

SpecT = TypeVar("SpecT")
ProviderT = TypeVar("ProviderT")


class PlatformEnum(enum.Enum):
    CUDA = enum.auto()
    ROCM = enum.auto()
    CPU = enum.auto()
    UNSPECIFIED = enum.auto()


@dataclass(frozen=True)
class DeviceCaps:
    platform: PlatformEnum
    device_type: str
    device_index: int
    device_name: str
    compute_capability: tuple[int, int] | None = None
    runtime_version: str | None = None
    supports_graph_capture: bool = False
    supports_torch_compile: bool = False
    supports_triton: bool = False
    supports_pin_memory: bool = False
    supports_bfloat16: bool = False
    supports_native_fp8: bool = False


def current_device_caps(device_index: int | None = None) -> DeviceCaps:
    """Caps of a CUDA device, or of the CPU when no GPU is visible (providers then refuse)."""
    if not torch.cuda.is_available():
        return DeviceCaps(PlatformEnum.CPU, "cpu", 0, "cpu")
    idx = torch.cuda.current_device() if device_index is None else int(device_index)
    props = torch.cuda.get_device_properties(idx)
    return DeviceCaps(PlatformEnum.CUDA, "cuda", idx, props.name, (props.major, props.minor),
                      torch.version.cuda, supports_graph_capture=True, supports_pin_memory=True,
                      supports_bfloat16=True, supports_native_fp8=props.major >= 9)


@dataclass(frozen=True)
class SupportResult:
    supported: bool
    reason: str

    @classmethod
    def yes(cls, reason: str = "supported") -> "SupportResult":
        return cls(True, reason)

    @classmethod
    def no(cls, reason: str) -> "SupportResult":
        return cls(False, reason)


_BINDINGS: dict[str, "weakref.WeakSet[object]"] = {}


def record_operator_binding(operator_type: str, provider: object) -> None:
    _BINDINGS.setdefault(operator_type, weakref.WeakSet()).add(provider)


def operator_runtime_stats() -> dict[str, list[dict[str, object]]]:
    """Per operator family: live providers grouped by name with their summed kernel-path counters."""
    out: dict[str, list[dict[str, object]]] = {}
    for family in sorted(_BINDINGS):
        by_name: dict[str, list[object]] = {}
        for prov in _BINDINGS[family]:
            by_name.setdefault(str(getattr(prov, "implementation_name", None) or prov.name), []).append(prov)
        rows = []
        for name in sorted(by_name):
            paths: dict[str, dict[str, int]] = {}
            fallbacks: dict[str, int] = {}
            instrumented = 0
            for prov in by_name[name]:
                fn = getattr(prov, "runtime_kernel_stats", None)
                if not callable(fn):
                    continue
                instrumented += 1
                stats = fn()
                for path, counts in stats.get("kernel_paths", {}).items():
                    agg = paths.setdefault(str(path), {})
                    for key, n in counts.items():
                        agg[str(key)] = agg.get(str(key), 0) + int(n)
                for reason, n in stats.get("fallback_reasons", {}).items():
                    fallbacks[str(reason)] = fallbacks.get(str(reason), 0) + int(n)
            rows.append({"implementation": name, "bound_provider_count": len(by_name[name]),
                         "instrumented_provider_count": instrumented,
                         "kernel_paths": {p: dict(sorted(c.items())) for p, c in sorted(paths.items())},
                         "fallback_reasons": dict(sorted(fallbacks.items()))})
        if rows:
            out[family] = rows
    return out


class OpRegistry(Generic[SpecT, ProviderT]):
    def __init__(self, family: str) -> None:
        self.family = str(family)
        self._providers: dict[str, type] = {}

    def register(self, provider: type) -> type:
        name = str(provider.name)
        if name in self._providers:
            raise ValueError(f"Provider {name!r} is already registered for {self.family!r}.")
        self._providers[name] = provider
        return provider

    @property
    def providers(self) -> tuple[type, ...]:
        return tuple(self._providers.values())


@dataclass(frozen=True)
class ResolvedProvider(Generic[ProviderT]):
    provider: ProviderT
    rejected: tuple[tuple[str, str], ...]


class OpResolver(Generic[SpecT, ProviderT]):
    def __init__(self, registry: OpRegistry) -> None:
        self.registry = registry

    def resolve(self, spec, caps: DeviceCaps, **provider_kwargs) -> ResolvedProvider:
        ok, rejected = [], []
        for prov in self.registry.providers:
            res = prov.supports(spec, caps)
            (ok if res.supported else rejected).append(prov if res.supported else (prov.name, res.reason))
        if not ok:
            details = "; ".join(f"{n}: {r}" for n, r in rejected)
            raise RuntimeError(f"No {self.registry.family} provider supports spec={spec!r} on "
                               f"device={caps.device_name!r}: {details or 'no providers registered'}.")
        ok.sort(key=lambda p: (-int(p.priority), p.name))
        chosen = ok[0](**provider_kwargs)
        record_operator_binding(self.registry.family, chosen)
        return ResolvedProvider(chosen, tuple(rejected))
