"""All-reduce operator family of the tensor-parallel decode path and its sm_100a provider.

Shape of the family follows the reference's ``operators/all_reduce.py`` (``AllReduceOpSpec`` 27-45,
``AllReduceProvider`` 48-73, ``TorchDistributedAllReduceProvider`` 412-429, ``PreparedAllReduceOp`` 459-541,
``prepare_all_reduce_op`` 518-541): the reference ships two FlashInfer providers gated to H100
(115-409) and the NCCL one; this package registers a one-shot all-reduce over NVLink peer memory written
for B200 (csrc/collective.cu) above the NCCL provider, which stays as the checked reference for it.
The call site is ``RowParallelLinear.forward`` (layers/linear.py:581-589): SUM of ``[B, hidden]`` over the
TP ranks after a row-parallel o_proj, once per layer, captured in the decode graph.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.distributed as dist

from .. import ops
from ..parallel import PeerComm
from .registry import DeviceCaps, OpRegistry, OpResolver, PlatformEnum, SupportResult, current_device_caps

ONESHOT_MAX_BYTES = 2 << 20  # above this a bandwidth-optimal (two-shot / NCCL) algorithm wins


@dataclass(frozen=True)
class AllReduceOpSpec:
    world_size: int
    ranks: tuple[int, ...]
    max_rows: int
    hidden_size: int
    dtype: torch.dtype
    cuda_graph: bool
    backend: str

    def __post_init__(self) -> None:
        if self.world_size <= 0 or self.max_rows <= 0 or self.hidden_size <= 0:
            raise ValueError("All-reduce dimensions must be positive.")
        if len(self.ranks) != self.world_size or len(set(self.ranks)) != self.world_size:
            raise ValueError(f"All-reduce ranks must contain world_size unique ranks, got {self.ranks}.")


class AllReduceProvider:
    name = ""
    priority = 0

    def prepare(self, spec: AllReduceOpSpec, *, group, rank: int, device_index: int | None = None) -> None:
        del spec, group, rank, device_index

    def close(self) -> None:
        pass

    def run(self, spec: AllReduceOpSpec, tensor: torch.Tensor, *, group) -> torch.Tensor:
        """Return the reduced tensor, which may or may not alias the input."""
        raise NotImplementedError


ALL_REDUCE_REGISTRY: OpRegistry = OpRegistry("all-reduce")


@ALL_REDUCE_REGISTRY.register
class B200NvlinkOneShotAllReduceProvider(AllReduceProvider):
    """One launch per all-reduce: every rank reads all inputs through NVLink peer mappings and sums them in
    rank order (svb200_allreduce_sum_bf16).  ``input_buffer()`` hands out the peer-mapped staging tensor the
    producer (o_proj) should write into, so that ``run`` moves no bytes of its own; a tensor that lives
    elsewhere is copied into it first."""

    name = "b200_nvlink_oneshot"
    priority = 200

    @classmethod
    def supports(cls, spec: AllReduceOpSpec, caps: DeviceCaps) -> SupportResult:
        if caps.platform != PlatformEnum.CUDA or caps.compute_capability != (10, 0):
            return SupportResult.no(f"requires CUDA SM100 (B200), got {caps.platform.name} {caps.compute_capability}")
        if spec.dtype != torch.bfloat16:
            return SupportResult.no(f"requires bf16, got {spec.dtype}")
        if not 2 <= spec.world_size <= 8:
            return SupportResult.no(f"serves 2..8 ranks of one NVLink domain, got {spec.world_size}")
        if spec.hidden_size % 8:
            return SupportResult.no(f"hidden_size={spec.hidden_size} must be a multiple of 8")
        if spec.max_rows * spec.hidden_size * 2 > ONESHOT_MAX_BYTES:
            return SupportResult.no("message too large for the one-shot algorithm")
        if spec.backend not in ("nccl", "gloo"):
            return SupportResult.no(f"needs an initialised process group to exchange IPC handles, backend={spec.backend!r}")
        return SupportResult.yes()

    def __init__(self) -> None:
        self.comm: PeerComm | None = None
        self._slot_bytes = 0
        self._next = 0
        self._out: torch.Tensor | None = None

    def prepare(self, spec, *, group, rank, device_index=None) -> None:
        del rank, device_index
        self._slot_bytes = (spec.max_rows * spec.hidden_size * 2 + 255) // 256 * 256
        # two staging slots: a rank may be one collective ahead of a peer that still reads the previous input
        self.comm = PeerComm(2 * self._slot_bytes, group=group)
        self._out = torch.empty((spec.max_rows, spec.hidden_size), dtype=torch.bfloat16, device=self.comm.device)

    def input_buffer(self, rows: int, hidden: int) -> torch.Tensor:
        """Staging tensor of the NEXT ``run`` (peer-mapped): write the partial sums here to skip the copy."""
        assert self.comm is not None, "prepare() first"
        return self.comm.view(self._next * self._slot_bytes, (rows, hidden), torch.bfloat16)

    def run(self, spec, tensor, *, group) -> torch.Tensor:
        del group
        assert self.comm is not None and self._out is not None, "prepare() first"
        rows = tensor.numel() // spec.hidden_size
        stage = self.input_buffer(rows, spec.hidden_size)
        if tensor.data_ptr() != stage.data_ptr():
            stage.copy_(tensor.view(rows, spec.hidden_size))
        out = self._out[:rows].view(tensor.shape)
        ops.allreduce_sum_bf16(self.comm, stage, out)
        self._next ^= 1
        return out

    def close(self) -> None:
        if self.comm is not None:
            self.comm.close()
            self.comm = None


@ALL_REDUCE_REGISTRY.register
class TorchDistributedAllReduceProvider(AllReduceProvider):
    """NCCL through ``torch.distributed`` (operators/all_reduce.py:412-429): the plumbing default and the
    reference the NVLink provider is checked against."""

    name = "torch_distributed"
    priority = 10

    @classmethod
    def supports(cls, spec: AllReduceOpSpec, caps: DeviceCaps) -> SupportResult:
        del spec, caps
        return SupportResult.yes()

    def run(self, spec, tensor, *, group) -> torch.Tensor:
        if spec.world_size > 1:
            if group is None:
                raise RuntimeError("Torch distributed all-reduce requires a process group when world_size > 1.")
            dist.all_reduce(tensor, group=group)
        return tensor


def _validate_tensor_contract(spec: AllReduceOpSpec, tensor: torch.Tensor) -> None:
    if tensor.dtype != spec.dtype:
        raise TypeError(f"All-reduce expected dtype={spec.dtype}, got {tensor.dtype}.")
    if tensor.ndim < 2:
        raise ValueError(f"All-reduce expects [..., hidden], got shape={tuple(tensor.shape)}.")
    row_count = tensor.numel() // int(tensor.shape[-1])
    if not 0 < row_count <= spec.max_rows:
        raise ValueError(f"All-reduce row count is outside the prepared range: rows={row_count} max_rows={spec.max_rows}.")
    if int(tensor.shape[-1]) != spec.hidden_size:
        raise ValueError(f"All-reduce hidden size does not match the prepared operator: hidden={int(tensor.shape[-1])} "
                         f"expected={spec.hidden_size}.")
    if not tensor.is_contiguous():
        raise ValueError(f"All-reduce requires a contiguous tensor, got stride={tensor.stride()}.")


class PreparedAllReduceOp:
    """A pre-bound all-reduce with no execution-time fallback."""

    def __init__(self, spec: AllReduceOpSpec, provider: AllReduceProvider, *, group) -> None:
        self.spec, self.provider, self.group = spec, provider, group
        self._closed = False

    @property
    def name(self) -> str:
        return self.provider.name

    def run(self, tensor: torch.Tensor) -> torch.Tensor:
        if self._closed:
            raise RuntimeError("All-reduce operator is closed.")
        _validate_tensor_contract(self.spec, tensor)
        output = self.provider.run(self.spec, tensor, group=self.group)
        _validate_tensor_contract(self.spec, output)
        if output.shape != tensor.shape or output.device != tensor.device:
            raise ValueError(f"All-reduce provider {self.provider.name} returned an incompatible tensor.")
        return output

    def close(self) -> None:
        if not self._closed:
            self.provider.close()
            self._closed = True


def prepare_all_reduce_op(spec: AllReduceOpSpec, *, group, rank: int, device_index: int | None = None,
                          provider: AllReduceProvider | None = None, caps: DeviceCaps | None = None) -> PreparedAllReduceOp:
    rank = int(rank)
    if not 0 <= rank < spec.world_size:
        raise ValueError(f"All-reduce rank must be in [0, {spec.world_size}), got {rank}.")
    if spec.world_size > 1 and group is None and not dist.is_initialized():
        raise RuntimeError("All-reduce requires a process group when world_size > 1.")
    if provider is None:
        if caps is None:
            caps = current_device_caps(device_index)
        provider = OpResolver(ALL_REDUCE_REGISTRY).resolve(spec, caps).provider
    provider.prepare(spec, group=group, rank=rank, device_index=device_index)
    return PreparedAllReduceOp(spec, provider, group=group)
