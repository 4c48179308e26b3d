"""Error behaviour of the host layer (CPU, 1-rank gloo group, checker backend): what is outside the
path raises with a clear message instead of silently doing something else."""
import os
import sys

import pytest
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)


@pytest.fixture(scope="module")
def pg(tmp_path_factory):
    store = tmp_path_factory.mktemp("pg") / "store"
    dist.init_process_group("gloo", init_method=f"file://{store}", rank=0, world_size=1)
    yield
    dist.destroy_process_group()


def _net():
    torch.manual_seed(0)
    return torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.ReLU(), torch.nn.Linear(16, 4))


def test_unsupported_constructor_options(native_lib, pg):
    import ddp_b200
    from oracle_backend import OracleKernels
    with pytest.raises(NotImplementedError, match="find_unused_parameters"):
        ddp_b200.DistributedDataParallel(_net(), find_unused_parameters=True, kernels=OracleKernels())
    with pytest.raises(RuntimeError, match="fused_allreduce"):
        ddp_b200.DistributedDataParallel(_net(), shard_optimizer=True, kernels=OracleKernels())
    frozen = _net()
    for p in frozen.parameters():
        p.requires_grad_(False)
    with pytest.raises(RuntimeError, match="doesn't have any"):
        ddp_b200.DistributedDataParallel(frozen, kernels=OracleKernels())
    emb = torch.nn.Sequential(torch.nn.Embedding(10, 4, sparse=True))
    with pytest.raises(NotImplementedError, match="sparse"):
        ddp_b200.DistributedDataParallel(emb, kernels=OracleKernels())
    with pytest.raises(NotImplementedError):
        ddp_b200.AdamW(_net().parameters(), differentiable=True)
    with pytest.raises(ValueError):
        ddp_b200.AdamW(_net().parameters(), lr=-1.0)
    with pytest.raises(ValueError):
        ddp_b200.AdamW(_net().parameters(), betas=(1.0, 0.999))


def test_unused_parameter_is_reported(native_lib, pg):
    """A parameter that gets no gradient: torch needs find_unused_parameters for that; here the end
    of backward says so instead of hanging or skipping the reduction."""
    import ddp_b200
    from oracle_backend import OracleKernels

    class Partial(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.used = torch.nn.Linear(4, 4)
            self.unused = torch.nn.Linear(4, 4)

        def forward(self, x):
            return self.used(x)

    ddp = ddp_b200.DistributedDataParallel(Partial(), kernels=OracleKernels())
    with pytest.raises(RuntimeError, match="received no gradient"):
        ddp(torch.randn(2, 4)).sum().backward()


def test_clip_and_step_guard_rails(native_lib, pg):
    import ddp_b200
    from oracle_backend import OracleKernels
    k = OracleKernels()
    net = _net()
    ddp = ddp_b200.DistributedDataParallel(net, kernels=k)
    with pytest.raises(RuntimeError, match="completed synchronised backward"):
        ddp.clip_grad_norm_(1.0)
    with pytest.raises(NotImplementedError):
        ddp_b200.clip_grad_norm_(net.parameters(), 1.0, norm_type=1.0, kernels=k)
    assert float(ddp_b200.clip_grad_norm_(net.parameters(), 1.0, kernels=k)) == 0.0      # no gradients yet
    ddp(torch.randn(2, 8)).sum().backward()
    total = ddp.clip_grad_norm_(1e9)
    want = torch.nn.utils.get_total_norm([p.grad for p in net.parameters()])
    assert abs(float(total) - float(want)) <= 1e-6 * float(want)
    # no_sync: nothing is reduced, and the overlapped norm is not available for that backward
    # (also when the previous clip kept the sums valid by handing the coefficient to the optimizer)
    ddp(torch.randn(2, 8)).sum().backward()
    ddp.clip_grad_norm_(1e9, fused_optimizer=ddp_b200.AdamW(net.parameters(), lr=1e-3, kernels=k))
    with ddp.no_sync():
        ddp(torch.randn(2, 8)).sum().backward()
    with pytest.raises(RuntimeError, match="completed synchronised backward"):
        ddp.clip_grad_norm_(1.0)
    # non-contiguous gradient is refused by the optimizer (the kernels stream flat memory)
    opt = ddp_b200.AdamW(net.parameters(), lr=1e-3, kernels=k)
    w = net[0].weight
    w.grad = torch.randn(8, 16).t()
    with pytest.raises(RuntimeError, match="contiguous"):
        opt.step()
