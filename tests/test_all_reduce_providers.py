"""All-reduce provider family (reference pattern: operators/all_reduce.py, tests/test_operator_providers.py:1089-1092):
resolution by device capabilities with fake DeviceCaps, the NCCL/gloo provider as the fallback everywhere else,
and the tensor contract of the prepared operator.  CPU only."""
from __future__ import annotations

import pytest
import torch

from sparse_vllm_b200.operators.all_reduce import (ALL_REDUCE_REGISTRY, AllReduceOpSpec,
                                                   B200NvlinkOneShotAllReduceProvider, PreparedAllReduceOp,
                                                   TorchDistributedAllReduceProvider, prepare_all_reduce_op)
from sparse_vllm_b200.operators.registry import DeviceCaps, OpResolver, PlatformEnum


def _caps(cc):
    return DeviceCaps(PlatformEnum.CUDA, "cuda", 0, "fake", cc, "12.9")


def _spec(**kw):
    base = dict(world_size=8, ranks=tuple(range(8)), max_rows=16, hidden_size=5120, dtype=torch.bfloat16, cuda_graph=True,
                backend="nccl")
    base.update(kw)
    return AllReduceOpSpec(**base)


def test_b200_resolves_the_nvlink_provider_and_everything_else_nccl():
    resolver = OpResolver(ALL_REDUCE_REGISTRY)
    assert isinstance(resolver.resolve(_spec(), _caps((10, 0))).provider, B200NvlinkOneShotAllReduceProvider)
    for spec, caps in ((_spec(), _caps((9, 0))),                                   # H100: not this provider's device
                       (_spec(dtype=torch.float16), _caps((10, 0))),               # bf16 only
                       (_spec(max_rows=4096, hidden_size=8192), _caps((10, 0))),   # too large for one shot
                       (_spec(world_size=1, ranks=(0,)), _caps((10, 0))),
                       (_spec(), DeviceCaps(PlatformEnum.CPU, "cpu", 0, "cpu"))):
        res = resolver.resolve(spec, caps)
        assert isinstance(res.provider, TorchDistributedAllReduceProvider)
        assert res.rejected and res.rejected[0][0] == "b200_nvlink_oneshot"


def test_spec_and_tensor_contract():
    with pytest.raises(ValueError, match="unique ranks"):
        AllReduceOpSpec(world_size=2, ranks=(0, 0), max_rows=1, hidden_size=8, dtype=torch.bfloat16, cuda_graph=False, backend="gloo")
    spec = _spec(world_size=1, ranks=(0,), backend="none")
    op = prepare_all_reduce_op(spec, group=None, rank=0, caps=DeviceCaps(PlatformEnum.CPU, "cpu", 0, "cpu"))
    assert isinstance(op, PreparedAllReduceOp) and op.name == "torch_distributed"
    x = torch.ones(3, 5120, dtype=torch.bfloat16)
    assert op.run(x) is x                                      # world size 1: identity, may alias the input
    with pytest.raises(TypeError):
        op.run(x.float())
    with pytest.raises(ValueError, match="row count"):
        op.run(torch.ones(17, 5120, dtype=torch.bfloat16))
    with pytest.raises(ValueError, match="hidden size"):
        op.run(torch.ones(3, 64, dtype=torch.bfloat16))
    op.close()
    with pytest.raises(RuntimeError, match="closed"):
        op.run(x)
