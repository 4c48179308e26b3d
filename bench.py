#!/usr/bin/env python
"""bench.py — the DDP gradient-sync + clip + AdamW hot path on B200.

A *step* is one pass of the hot path over one set of synthetic gradients for the Llama-3.2-1B
architecture in bf16 (BASELINE.json configs[1]: 146 parameter tensors, 1,235,814,400 elements).
N=1: nothing to exchange, so one coalesced sum-of-squares launch over the gradients where they lie, the total
norm + clip coefficient on the device, ONE fused clip+AdamW launch over all parameters. N>1: the gradients live
in the bucket arena (gradient_as_bucket_view), so a step is the fused allreduce+norm kernel over NVLink peer
memory / NVSwitch multicast with the 1/world_size scale folded in (or NCCL + a norm launch), finalize, AdamW;
--separate-grads adds the pack pass. Before the timed loop an N>1 run checks its collective against NCCL, across
ranks and against the sharded update (`parity_check`; a failure is exit code 3). Forward/backward of the model
are not part of the path (the next row) — the N=1 line carries them as the `train` side record,
and torch's own CUDA kernels on the same tensors as `incumbent`.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--zero1] [--workload W]

N>1 is launched by torchrun (one rank per GPU, NCCL). Prints ONE JSON line on rank 0.
--impl reference times the reference's own CPU implementation of the same path: the UNMODIFIED
stock torch code (DistributedDataParallel on gloo + clip_grad_norm_ + torch.optim.AdamW) on the
box's host cores, on the full parameter set, fed the same kind of synthetic gradients through autograd.
"""
from __future__ import annotations

import argparse
import json
import os
import socket
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# stdout carries ONE JSON line: NCCL's own output (the version banner under NCCL_DEBUG=VERSION, warnings) goes to stderr.
# Set before torch loads the library.
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")

import torch
import torch.distributed as dist

METRIC = "ddp_hot_path_steps_per_s"
UNIT = "steps/s x ranks (one step = grad-sync + clip + AdamW over all parameters of one replica)"
MAX_NORM = 1.0
LR = 5e-5


# ---- workload --------------------------------------------------------------------------------
def llama_param_shapes(hidden, layers, heads, kv_heads, inter, vocab, tie):
    """Parameter shapes of transformers' LlamaForCausalLM in DDP registration order
    (tests/test_workload_shapes.py checks them against the golden numel lists)."""
    hd = hidden // heads
    shapes = [(vocab, hidden)]
    for _ in range(layers):
        shapes += [(hidden, hidden), (kv_heads * hd, hidden), (kv_heads * hd, hidden), (hidden, hidden),
                   (inter, hidden), (inter, hidden), (hidden, inter), (hidden,), (hidden,)]
    shapes.append((hidden,))
    if not tie:
        shapes.append((vocab, hidden))
    return shapes


WORKLOADS = {
    "llama-3.2-1b": dict(hidden=2048, layers=16, heads=32, kv_heads=8, inter=8192, vocab=128256, tie=True),
    "llama-3.1-8b": dict(hidden=4096, layers=32, heads=32, kv_heads=8, inter=14336, vocab=128256, tie=False),
    "llama-3.1-70b-4layer": dict(hidden=8192, layers=4, heads=64, kv_heads=8, inter=28672, vocab=128256, tie=False),
    "tiny": dict(hidden=256, layers=2, heads=8, kv_heads=2, inter=512, vocab=1024, tie=True),
}


def grad_ready_order(n_params, layers, tie):
    """Order in which a Llama backward makes gradients ready (SURVEY.md §8a probe): [lm_head,] final
    norm, then per layer from the last: down, up, gate, post_attn_ln, o, v, k, q, input_ln; embed last."""
    order = []
    if not tie:
        order.append(n_params - 1)
    order.append(1 + 9 * layers)
    for l in range(layers - 1, -1, -1):
        b = 1 + 9 * l
        order += [b + 6, b + 5, b + 4, b + 8, b + 3, b + 2, b + 1, b + 0, b + 7]
    order.append(0)
    return order


class ParamSet(torch.nn.Module):
    """The parameter tensors of the model, without its forward: the hot path only needs them."""

    def __init__(self, shapes, dtype, device, seed=0):
        super().__init__()
        g = torch.Generator(device=device).manual_seed(seed)
        self.ps = torch.nn.ParameterList(
            [torch.nn.Parameter(torch.empty(s, dtype=dtype, device=device).normal_(0, 0.02, generator=g))
             for s in shapes])


class SynthGrad(torch.autograd.Function):
    """loss = SynthGrad(*params): backward hands each parameter its preset synthetic gradient, so the
    stock autograd -> AccumulateGrad -> DDP-hook machinery runs exactly as in training."""

    @staticmethod
    def forward(ctx, holder, *params):
        ctx.holder = holder
        return params[0].new_zeros(())

    @staticmethod
    def backward(ctx, grad_out):
        return (None,) + tuple(ctx.holder["grads"])


# ---- helpers ----------------------------------------------------------------------------------
def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.proc, self.path = gpu_index, None, f"/tmp/ddpb200_clocks_{os.getpid()}.csv"

    def start(self):
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "25"], stdout=self.f,
                                         stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.f.close()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                mx.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(names, parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        try:
            os.remove(self.path)
        except OSError:
            pass
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


def measured_peaks():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(d["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (torch copy_, read+write)"
    except Exception:
        return 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"


def adamw_traffic_from_profile():
    """dram bytes per launch of the fused clip+AdamW kernel from the committed ncu summary, if any."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "adamw_ncu_summary.json")))
        return d.get("dram_bytes_per_launch_llama_1b")
    except Exception:
        return None


def bind_to_gpu_numa_node(dev_index):
    """Run this rank on the CPUs NVML lists as local to its GPU, so that the pinned host buffers of the
    end-to-end leg are first-touched on the GPU's NUMA node (8 ranks otherwise share one node's memory
    and root complex). Returns the number of CPUs bound to, or None if NVML is unavailable."""
    try:
        import pynvml
        pynvml.nvmlInit()
        uuid = str(torch.cuda.get_device_properties(dev_index).uuid)
        try:
            h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode())
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(dev_index)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = [w * 64 + b for w, word in enumerate(mask) for b in range(64) if (int(word) >> b) & 1]
        allowed = os.sched_getaffinity(0)
        cpus = [c for c in cpus if c in allowed]
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


# ---- N>1 self-check: the collective the step uses, against NCCL and across ranks -----------------------
def _ulp_key(t):
    """Monotonic integer key of a bf16/fp32 tensor's bit patterns (difference of keys = distance in ulps)."""
    if t.dtype == torch.bfloat16:
        b = t.view(torch.int16).to(torch.int32)
        return torch.where(b < 0, -(b & 0x7fff), b)
    b = t.view(torch.int32).to(torch.int64)
    return torch.where(b < 0, -(b & 0x7fffffff), b)


def _max_ulp(a, b, block=1 << 27):
    worst = 0
    for o in range(0, a.numel(), block):
        worst = max(worst, int((_ulp_key(a[o:o + block]) - _ulp_key(b[o:o + block])).abs().max()))
    return worst


def _ulp_of(x, dt):
    """Spacing of `dt` at |x| (fp32 tensor in, fp32 out); the smallest subnormal step at 0."""
    mant = 8 if dt == torch.bfloat16 else 24
    _, e = torch.frexp(x.abs())                                  # |x| = m * 2^e, m in [0.5, 1)
    e = torch.where(x == 0, torch.full_like(e, -200), e)
    return torch.ldexp(torch.ones_like(x), (e - mant).clamp_min(-126 - mant + 1))


def _err_in_ulps(got, want32, sumabs32, dt, block=1 << 26):
    """max over elements of (|got - want| - fp32 noise) / ulp_dt(max(|got|, |want|)), where the noise allowance is
    2^-22 * (sum over ranks of |input|): the two fp32 sums being compared add in different orders. Under
    cancellation the result is small against its inputs and a plain ulp distance would blame the rounding
    of the RESULT for the accumulation error of the INPUTS."""
    worst = 0.0
    for o in range(0, got.numel(), block):
        g, w = got[o:o + block].float(), want32[o:o + block]
        excess = ((g - w).abs() - sumabs32[o:o + block] * 2.0 ** -22).clamp_min(0)
        worst = max(worst, float((excess / _ulp_of(torch.maximum(g.abs(), w.abs()), dt)).max()))
    return worst


def parity_check(ddp, ws, rank, dev, zero1_too, shapes, order, cfg):
    """Run BEFORE the timed loop on the real bucket arena of this run (multi-GPU gradients
    are checked to a tolerance, everything replicated must be bit-identical):
      (i)   the step's gradient collective (fused kernel with the scale it is given in the step, or NCCL)
            against stock DDP's arithmetic — round(fp32(g) * float(1/ws)) per rank, then SUM — with the sum
            taken by NCCL in fp32: within one rounding of the bucket dtype; and against NCCL's own
            allreduce in the bucket dtype (reported, NCCL rounds at every hop);
      (ii)  averaged gradients and per-tensor sums of squares bit-identical on all ranks;
      (iii) sums of squares against a double-precision sum of what was stored, 3e-6 relative;
      (iv)  sharded update (ZeRO-1 style): parameters after 2 steps bit-equal to the replicated path.
    Every rank returns the same verdict."""
    k = ddp._kernels
    inv = ddp._inv_world_size
    out = {"world": ws, "collective": "fused" if ddp._fused else "nccl"}
    ok = True
    worst_f32, worst_nccl, worst_rel, same, worst_err, nccl_err, differ_nccl = 0, 0, 0.0, True, 0.0, 0.0, 0
    arenas = ddp._sym if ddp._fused else {dt: None for dt in ddp._arenas}
    for dt in arenas:
        flat = ddp._arenas[dt]
        rows = [bk for bk in ddp._buckets if bk.dtype == dt]
        first, n = rows[0].first, sum(bk.n for bk in rows)
        flat.normal_(0, 0.01, generator=torch.Generator(device=dev).manual_seed(1234 + rank))
        pre = (flat.float() * inv).to(dt)                        # stock: bucket_view = grad * (1/ws), in the bucket dtype
        ref32 = pre.float()
        dist.all_reduce(ref32)                                   # SUM of the pre-scaled values, fp32 accumulation
        sumabs = pre.float().abs_()
        dist.all_reduce(sumabs)
        want = ref32.to(dt)
        nccl = pre.clone()
        dist.all_reduce(nccl)                                    # NCCL in the bucket dtype
        ss = torch.empty(n, dtype=torch.float32, device=dev)
        torch.cuda.synchronize()
        if ddp._fused:
            fold = ddp._folds_scale(dt)
            if not fold:
                flat.copy_(pre)
            k.allreduce_sumsq(ddp._view_table, dt, ddp._sym[dt], ss, first=first, count=n, scatter=False,
                              scale=inv if fold else 1.0)
            out["scale_folded_into_collective"] = bool(fold)
        else:
            flat.copy_(pre)
            dist.all_reduce(flat)
            k.sumsq(ddp._view_table, dt, ss, first=first, count=n)
        del pre
        torch.cuda.synchronize()
        worst_f32 = max(worst_f32, _max_ulp(flat, want))
        worst_nccl = max(worst_nccl, _max_ulp(flat, nccl))
        worst_err = max(worst_err, _err_in_ulps(flat, ref32, sumabs, dt))
        nccl_err = max(nccl_err, _err_in_ulps(nccl, ref32, sumabs, dt))
        differ_nccl += int((flat != nccl).sum())
        del want, nccl, ref32, sumabs
        views = [v for bk in rows for v in bk.views]
        exact = torch.stack([v.double().pow(2).sum() for v in views])
        rel = ((ss.double() - exact).abs() / exact.clamp_min(1e-300)).max()
        worst_rel = max(worst_rel, float(rel))
        twin = flat.clone()
        dist.broadcast(twin, src=0)
        same = same and torch.equal(twin.view(torch.int16 if dt == torch.bfloat16 else torch.int32),
                                    flat.view(torch.int16 if dt == torch.bfloat16 else torch.int32))
        del twin
        ss0 = ss.clone()
        dist.broadcast(ss0, src=0)
        same = same and torch.equal(ss0.view(torch.int32), ss.view(torch.int32))
    flag = torch.tensor([1 if same else 0, worst_f32, worst_nccl], dtype=torch.int64, device=dev)
    dist.all_reduce(flag[:1], op=dist.ReduceOp.MIN)
    dist.all_reduce(flag[1:], op=dist.ReduceOp.MAX)
    relt = torch.tensor([worst_rel, worst_err, nccl_err], dtype=torch.float64, device=dev)
    dist.all_reduce(relt, op=dist.ReduceOp.MAX)
    name = "fused" if ddp._fused else "nccl"
    mc = ddp._fused and any(a.use_multicast for a in ddp._sym.values())
    # |result - fp32 sum| in ulps of the bucket dtype at the result, after an allowance for fp32 accumulation
    # order (see _err_in_ulps). Peer mode: one owner adds in fp32 and rounds once -> <= 0.5. Multicast: the
    # NVSwitch's adder is up to one ulp off round-to-nearest (tools/mc_probe.py) -> <= 1.5. NCCL in the bucket
    # dtype may round at every hop -> <= world.
    limit = (1.5 if mc else 0.5) if ddp._fused else float(ws)
    out[f"{name}_vs_fp32sum_err_ulps"] = float(relt[1])
    out["err_ulps_limit"] = limit * 1.001
    out["nccl_bucket_dtype_vs_fp32sum_err_ulps"] = float(relt[2])
    out[f"{name}_vs_fp32sum_max_ulp_distance"] = int(flag[1])         # plain ulp distances, for the record: they
    out[f"{name}_vs_nccl_max_ulp_distance"] = int(flag[2])            # blow up where the ranks' values cancel
    out[f"{name}_vs_nccl_elements_differing_rank0"] = differ_nccl
    out["transport"] = ("multicast (NVSwitch adds)" if mc else "peer (owner adds in fp32, rank order)") if ddp._fused else "nccl"
    out["ranks_bit_identical"] = bool(int(flag[0]))
    out["sumsq_rel"] = float(relt[0])
    ok = ok and out["ranks_bit_identical"] and out["sumsq_rel"] <= 3e-6 and float(relt[1]) <= limit * 1.001
    if zero1_too:
        try:
            z = zero1_parity(shapes, order, dev, rank, ws)
        except Exception as e:                                   # symmetric memory could not be set up: say so
            z = {"zero1_bit_identical": None, "zero1_note": f"{type(e).__name__}: {str(e)[:160]}"}
        out.update(z)
        if z.get("zero1_bit_identical") is False:
            ok = False
    out["ok"] = bool(ok)
    return out


def zero1_parity(shapes, order, dev, rank, ws, steps=2):
    """(iv): the sharded update next to the replicated path on twin copies of the full parameter set, same
    rank-dependent synthetic gradients, both over the peer-memory transport (so that both sum in rank
    order): parameters after `steps` steps must be the same bits."""
    import ddp_b200
    local = dev.index
    dtype = torch.bfloat16
    prev_mode = os.environ.get("DDPB200_FUSED_MODE")
    os.environ["DDPB200_FUSED_MODE"] = "peer"       # read when an arena is created (also by the bucket rebuild below)
    try:
        return _zero1_parity(ddp_b200, shapes, order, dev, rank, local, dtype, steps)
    finally:
        if prev_mode is None:
            os.environ.pop("DDPB200_FUSED_MODE", None)
        else:
            os.environ["DDPB200_FUSED_MODE"] = prev_mode


def _zero1_parity(ddp_b200, shapes, order, dev, rank, local, dtype, steps):
    mA, mB = ParamSet(shapes, dtype, dev, seed=3), ParamSet(shapes, dtype, dev, seed=3)
    dA = ddp_b200.DistributedDataParallel(mA, device_ids=[local], gradient_as_bucket_view=True, fused_allreduce=True)
    dB = ddp_b200.DistributedDataParallel(mB, device_ids=[local], gradient_as_bucket_view=True, fused_allreduce=True,
                                          shard_optimizer=True)
    oA = ddp_b200.AdamW(list(mA.parameters()), lr=LR, weight_decay=1e-2)
    oB = ddp_b200.ShardedAdamW(dB, lr=LR, weight_decay=1e-2)
    pA, pB = list(mA.parameters()), list(mB.parameters())
    numel = [p.numel() for p in pA]
    total = sum(numel)
    same = True
    for s in range(steps):
        gflat = torch.empty(total, dtype=dtype, device=dev).normal_(
            0, 0.01, generator=torch.Generator(device=dev).manual_seed(100 * s + 7 + rank))
        gs, o = [], 0
        for n_, sh in zip(numel, shapes):
            gs.append(gflat[o:o + n_].view(sh))
            o += n_
        dA.reduce_gradients([g.clone() for g in gs], order=order if s == 0 else None)
        dA.clip_grad_norm_(MAX_NORM, fused_optimizer=oA)
        oA.step()
        dB.reduce_gradients(gs)
        dB.clip_grad_norm_(MAX_NORM, fused_optimizer=oB)
        oB.step()
        torch.cuda.synchronize()
        same = same and all(torch.equal(a.view(torch.int16), b.view(torch.int16)) for a, b in zip(pA, pB))
    changed = not torch.equal(pA[0], ParamSet(shapes[:1], dtype, dev, seed=3).ps[0])
    flag = torch.tensor([1 if (same and changed) else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    del dA, dB, oA, oB, mA, mB
    torch.cuda.empty_cache()
    return {"zero1_bit_identical": bool(int(flag)), "zero1_steps": steps, "zero1_transport": "peer (both arms)"}



# ---- N=1 side records: the whole training step, and torch's own CUDA kernels on the same tensors ----------
def train_record(dev, steps=8, warmup=3, batch=4, seq=2048, model_name="llama-1b"):
    """the step either side of the hot path. forward + backward of the random-init
    Llama-3.2-1B-architecture LM (REF/src/distributed_inference.py:61-69 is where the loop lives,
    REF/src/utils.py:12-23 the seam) on `batch` x `seq` synthetic tokens from pinned host memory, loss read
    back, with the gradient sync + clip + update done by ddp_b200 and by stock torch
    (DistributedDataParallel + clip_grad_norm_ + AdamW(fused=True)) on the same GPU, weights and batches.
    Two precisions: "bf16" = the whole model cast to bf16 (p/g/m/v bf16); "autocast" = fp32 master
    weights, forward under torch.autocast(bfloat16), fp32 gradients and optimizer state."""
    import copy
    sys.path.insert(0, os.path.join(ROOT, "src"))
    import ddp_b200
    from distributed_finetuning import build_model
    out = {"model": model_name + " architecture, random init", "tokens_per_step": batch * seq, "batch": batch, "seq": seq,
           "steps": steps, "warmup": warmup, "data": "synthetic token ids from pinned host memory; loss read back every step"}
    local = dev.index
    for mode in ("bf16", "autocast"):
        pdt = torch.bfloat16 if mode == "bf16" else torch.float32
        base = build_model(model_name, dev, pdt)
        vocab = base.config.vocab_size
        g = torch.Generator().manual_seed(1000)
        n_steps = steps if mode == "bf16" else max(3, steps // 2)
        host_ids = [torch.randint(0, vocab, (batch, seq), generator=g).pin_memory() for _ in range(n_steps + warmup)]
        loss_host = torch.empty((), dtype=torch.float32).pin_memory()
        rec = {}
        for impl in ("ours", "stock"):
            model = copy.deepcopy(base)
            if impl == "ours":
                ddp = ddp_b200.DistributedDataParallel(model, device_ids=[local], gradient_as_bucket_view=True)
                opt = ddp_b200.AdamW(model.parameters(), lr=LR)
                clip = lambda: ddp.clip_grad_norm_(MAX_NORM, fused_optimizer=opt)                  # noqa: E731
            else:
                ddp = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local], gradient_as_bucket_view=True)
                opt = torch.optim.AdamW(model.parameters(), lr=LR, fused=True)
                clip = lambda: torch.nn.utils.clip_grad_norm_(model.parameters(), MAX_NORM)        # noqa: E731
            tail0 = [torch.cuda.Event(enable_timing=True) for _ in range(n_steps)]
            tail1 = [torch.cuda.Event(enable_timing=True) for _ in range(n_steps)]

            def one(i, k=None):
                ids = host_ids[i].to(dev, non_blocking=True)
                with torch.autocast("cuda", dtype=torch.bfloat16, enabled=(mode == "autocast")):
                    loss = ddp(input_ids=ids, labels=ids).loss
                loss.backward()
                if k is not None:
                    tail0[k].record()          # from here: what is left of sync + clip + update after backward
                clip()
                opt.step()
                if k is not None:
                    tail1[k].record()
                opt.zero_grad()
                loss_host.copy_(loss.detach().float(), non_blocking=True)

            for i in range(warmup):
                one(i)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for k in range(n_steps):
                one(warmup + k, k)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / n_steps
            rec[impl] = {"ms_per_step": ms, "tokens_per_s": batch * seq * 1e3 / ms,
                         "ms_after_backward_clip_plus_update": sum(a.elapsed_time(b) for a, b in zip(tail0, tail1)) / n_steps,
                         "loss": float(loss_host), "steps": n_steps}
            del ddp, opt, model
            torch.cuda.empty_cache()
        rec["speedup_whole_step"] = rec["stock"]["ms_per_step"] / rec["ours"]["ms_per_step"]
        rec["speedup_after_backward"] = (rec["stock"]["ms_after_backward_clip_plus_update"] /
                                         rec["ours"]["ms_after_backward_clip_plus_update"])
        out[mode] = rec
        del base
        torch.cuda.empty_cache()
    out["ours_tokens_per_s"] = out["bf16"]["ours"]["tokens_per_s"]
    out["stock_tokens_per_s"] = out["bf16"]["stock"]["tokens_per_s"]
    return out


def incumbent_record(params, grads, dev, reps=10):
    """torch's own CUDA kernels on the tensors of this bench: `_fused_adamw_` (torch.optim.AdamW(fused=True),
    TORCH/include/ATen/native/cuda/fused_adam_utils.cuh:108-214) and `_foreach_norm` + clip_grad_norm_
    (TORCH/nn/utils/clip_grad.py:96) against this library's kernels, CUDA events, same box, same run."""
    import ddp_b200
    K = ddp_b200.default_kernels()

    def timeit(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    for p, g in zip(params, grads):
        p.grad = g
    numel = [p.numel() for p in params]
    dt = params[0].dtype
    ss = torch.empty(len(params), dtype=torch.float32, device=dev)
    tab = K.table([None, grads, None, None], numel, device=dev)
    ours_norm = timeit(lambda: K.sumsq(tab, dt, ss))
    torch_norm = timeit(lambda: torch._foreach_norm(grads, 2))
    ours_opt = ddp_b200.AdamW(params, lr=LR, weight_decay=1e-2)

    def ours_tail():
        ddp_b200.clip_grad_norm_(params, MAX_NORM, fused_optimizer=ours_opt)
        ours_opt.step()
    ours_tail_ms = timeit(ours_tail)
    ours_adamw = timeit(ours_opt.step)
    del ours_opt
    torch.cuda.empty_cache()
    topt = torch.optim.AdamW(params, lr=LR, weight_decay=1e-2, fused=True)
    torch_adamw = timeit(topt.step)

    def torch_tail():
        torch.nn.utils.clip_grad_norm_(params, MAX_NORM)
        topt.step()
    torch_tail_ms = timeit(torch_tail)
    del topt
    for p in params:
        p.grad = None
    torch.cuda.empty_cache()
    return {"tensors": len(params), "elements": sum(numel), "dtype": str(dt).replace("torch.", ""), "reps": reps,
            "adamw_ms": {"ours": ours_adamw, "torch_fused_adamw_": torch_adamw},
            "adamw_speedup": torch_adamw / ours_adamw,
            "norm_ms": {"ours": ours_norm, "torch_foreach_norm": torch_norm}, "norm_speedup": torch_norm / ours_norm,
            "clip_plus_adamw_ms": {"ours": ours_tail_ms, "torch_clip_grad_norm_+fused_adamw": torch_tail_ms},
            "clip_plus_adamw_speedup": torch_tail_ms / ours_tail_ms}



# ---- our arm ------------------------------------------------------------------------------------
def run_ours(args):
    import ddp_b200

    rank = int(os.environ.get("RANK", 0))
    ws = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    assert ws == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={ws} (launch N>1 with torchrun)"
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    all_cpus = os.sched_getaffinity(0)
    numa_cpus = bind_to_gpu_numa_node(local) if not args.no_numa_bind else None
    if "MASTER_ADDR" not in os.environ:
        os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(free_port())
    dist.init_process_group("nccl", rank=rank, world_size=ws, device_id=dev)

    cfg = WORKLOADS[args.workload]
    shapes = llama_param_shapes(**cfg)
    dtype = torch.bfloat16
    model = ParamSet(shapes, dtype, dev)
    params = list(model.parameters())
    numel = [p.numel() for p in params]
    total = sum(numel)
    esize = 2
    fused_note = None
    zero1 = bool(args.zero1) and ws > 1
    try:
        ddp = ddp_b200.DistributedDataParallel(model, device_ids=[local], gradient_as_bucket_view=True,
                                               fused_allreduce=(args.allreduce == "fused" or zero1),
                                               shard_optimizer=zero1)
    except Exception as e:          # symmetric memory could not be set up on this box: say so, use NCCL
        if args.allreduce != "fused" or zero1:
            raise
        fused_note = f"fused allreduce unavailable ({type(e).__name__}: {str(e)[:120]}); NCCL used"
        print(fused_note, file=sys.stderr, flush=True)
        ddp = ddp_b200.DistributedDataParallel(model, device_ids=[local], gradient_as_bucket_view=True,
                                               fused_allreduce=False)
    if ws > 1 and args.allreduce == "fused" and not zero1:
        ok = torch.tensor([1 if ddp._fused else 0], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok) == 0 and ddp._fused:         # another rank fell back: all ranks must issue the same collectives
            fused_note = "fused allreduce unavailable on another rank; NCCL used"
            ddp = ddp_b200.DistributedDataParallel(model, device_ids=[local], gradient_as_bucket_view=True,
                                                   fused_allreduce=False)
    if zero1:
        opt = ddp_b200.ShardedAdamW(ddp, lr=LR, weight_decay=1e-2)
    else:
        opt = ddp_b200.AdamW(params, lr=LR, weight_decay=1e-2)

    # synthetic gradients, resident in HBM, one tensor per parameter (rank-dependent values)
    gflat = torch.empty(total, dtype=dtype, device=dev).normal_(0, 0.01, generator=torch.Generator(device=dev).manual_seed(1 + rank))
    offs = [0]
    for n in numel[:-1]:
        offs.append(offs[-1] + n)
    grads = [gflat[o:o + n].view(s) for o, n, s in zip(offs, numel, shapes)]
    order = grad_ready_order(len(params), cfg["layers"], cfg["tie"])

    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(args.steps)]

    # N>1: after the rebuild the gradients LIVE in the bucket arena (gradient_as_bucket_view: p.grad is its
    # bucket view, the producer writes there), so no pack launch is left and the 1/world_size scale rides in
    # the collective. --separate-grads keeps them in a buffer of their own (pack = copy + scale per step).
    in_arena = ws > 1 and not args.separate_grads
    state = {"resident": False}

    def step(i=None, first=False):
        if state["resident"]:
            ddp.reduce_gradients(None)                 # p.grad already is the bucket view
        else:
            ddp.reduce_gradients(grads, order=order if first else None)
        ddp.clip_grad_norm_(MAX_NORM, fused_optimizer=opt)
        if i is not None:
            ev[i][0].record()
        opt.step()
        if i is not None:
            ev[i][1].record()
        if in_arena and ddp._has_rebuilt and not first:
            state["resident"] = True                   # the arena keeps the (averaged) gradients: next step's input
        else:
            opt.zero_grad()

    def sync_all():
        torch.cuda.synchronize()
        if ws > 1:
            dist.barrier()
        torch.cuda.synchronize()

    step(first=True)                    # iteration 0: one bucket, records the ready order
    step()                              # rebuilt buckets from here on
    parity = None
    if ws > 1 and not args.no_parity_check:
        parity = parity_check(ddp, ws, rank, dev, zero1_too=not args.no_zero1_check, shapes=shapes,
                              order=order, cfg=cfg)
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()                  # samples the same load from the warm-up on, through the timed region
    for _ in range(max(args.warmup, 3)):
        step()
    stats0 = dict(ddp._get_ddp_logging_data())

    # ---- timed: inputs resident in HBM -----------------------------------------------------------
    sync_all()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for i in range(args.steps):
        step(i)
    t1.record()
    sync_all()
    clk = clocks.stop() if rank == 0 else None
    ms = t0.elapsed_time(t1)
    adamw_ms = statistics.mean(a.elapsed_time(b) for a, b in ev)
    stats1 = dict(ddp._get_ddp_logging_data())
    launches = sum(stats1[k] - stats0[k] for k in ("pack_launches", "norm_launches", "unpack_launches"))
    launches += args.steps * (1 + opt.launches_last_step)          # finalize + fused AdamW per step
    launches += 2 * (stats1.get("fused_allreduces", 0) - stats0.get("fused_allreduces", 0))   # allreduce kernel + fold
    allreduces = stats1["allreduces"] - stats0["allreduces"]

    # ---- side figure at N>1: the same step with the buckets launched ONE BY ONE, as the autograd hooks of a real
    # backward launch them (a collective per 25 MiB-capped bucket instead of one over the whole arena); in training
    # this path hides behind backward (the `train` record and tools/train_bench.py), here it stands alone
    per_bucket_ms = None
    if ws > 1:
        pb_steps = 5

        def step_per_bucket():
            ddp.reduce_gradients(None if state["resident"] else grads, order=list(ddp._order))
            ddp.clip_grad_norm_(MAX_NORM, fused_optimizer=opt)
            opt.step()
            if not state["resident"]:
                opt.zero_grad()
        step_per_bucket()
        sync_all()
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        p0.record()
        for _ in range(pb_steps):
            step_per_bucket()
        p1.record()
        sync_all()
        per_bucket_ms = p0.elapsed_time(p1) / pb_steps

    def run_e2e(e2e_steps):
        # ---- timed: end to end from HOST gradients ---------------------------------------------------
        # Each step's gradients arrive from pinned host memory (H2D inside the timed region) and the
        # pre-clip total norm is read back on the host (D2H). The copies run bucket by bucket on a copy
        # stream into one of two landing buffers, so a bucket is processed as soon as it has landed and
        # the update of step k overlaps the transfer of step k+1.
        layout = ddp.bucket_layout()
        order_now = [pi for b in layout for pi in b["parameters"]]
        host_g = torch.empty(total, dtype=dtype).pin_memory()
        host_g.copy_(gflat)
        land = [gflat, torch.empty_like(gflat)]
        land_views, bucket_span = [], []
        for buf in land:                         # landing buffers are laid out in bucket order
            views, o = [None] * len(params), 0
            for pi in order_now:
                views[pi] = buf[o:o + numel[pi]].view(shapes[pi])
                o += numel[pi]
            land_views.append(views)
        o = 0
        for b in layout:
            bucket_span.append((o, o + b["numel"]))
            o += b["numel"]
        copy_stream = torch.cuda.Stream(device=dev)
        landed = [[torch.cuda.Event() for _ in layout] for _ in land]
        consumed = [torch.cuda.Event() for _ in land]
        norm_ready = [torch.cuda.Event() for _ in land]
        norm_host2 = [torch.empty(2, dtype=torch.float32).pin_memory() for _ in land]
        for ev_ in consumed:
            ev_.record()

        def h2d(slot):
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(consumed[slot])             # the buffer's previous reader is done
                for b, (lo, hi) in enumerate(bucket_span):
                    land[slot][lo:hi].copy_(host_g[lo:hi], non_blocking=True)
                    landed[slot][b].record()

        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        last_norm = 0.0
        h2d(0)
        for it in range(e2e_steps):
            slot = it % 2
            if it + 1 < e2e_steps:
                h2d(1 - slot)                                      # next step's gradients start moving now
            cur = torch.cuda.current_stream()
            state["resident"] = False
            ddp.reduce_gradients(land_views[slot], before_bucket=lambda b, _idx: cur.wait_event(landed[slot][b]))
            ddp.clip_grad_norm_(MAX_NORM, fused_optimizer=opt)
            norm_host2[slot].copy_(ddp._norm_out, non_blocking=True)
            norm_ready[slot].record()
            opt.step()
            consumed[slot].record()
            opt.zero_grad()
            norm_ready[slot].synchronize()                         # the step's result, on the host
            last_norm = float(norm_host2[slot][0])
        e1.record()
        sync_all()
        e2e_ms = e0.elapsed_time(e1)
        return e2e_ms, last_norm

    e2e_steps = 0 if args.skip_e2e else max(2, min(args.steps, 10))
    if e2e_steps:
        e2e_ms, last_norm = run_e2e(e2e_steps)
    else:
        e2e_ms, last_norm = float("nan"), float(ddp._norm_out[0])

    t = torch.tensor([ms, e2e_ms, adamw_ms, per_bucket_ms or 0.0], dtype=torch.float64, device=dev)
    if ws > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms, adamw_ms, per_bucket_ms = [float(x) for x in t.tolist()]

    out = None
    if rank == 0:
        peak, peak_src = measured_peaks()
        alg_bytes = 14 * total // (ws if zero1 else 1)                # bf16 p,g,m,v read + p,m,v write (1/ws of them if sharded)
        achieved = alg_bytes / (adamw_ms * 1e-3) / 1e9
        bucket_bytes = stats1["bucket_sizes"]
        out = {
            "metric": METRIC, "value": ws * args.steps / (ms * 1e-3), "unit": UNIT, "n_gpus": ws,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "arithmetic": "fp32 opmath on bf16 storage (p, g, exp_avg, exp_avg_sq), as torch's fused AdamW",
            "data": "synthetic", "impl": "ours" + ("-zero1" if zero1 else ""),
            "config": workload_config(args.workload, len(params), total, ws),
            "implementation": f"ddp_b200: gradient_as_bucket_view, {len(bucket_bytes)} buckets after rebuild" +
                              (", sharded optimizer (ZeRO-1 style)" if zero1 else ""),
            "steps_per_s": args.steps / (ms * 1e-3),
            "e2e": {"value": (ws * e2e_steps / (e2e_ms * 1e-3)) if e2e_steps else None, "unit": UNIT, "steps": e2e_steps,
                    "h2d_bytes_per_step": total * esize, "d2h_bytes_per_step": 8, "numa_local_cpus": numa_cpus,
                    "note": "gradients of each step copied from pinned host memory (bucket by bucket, double "
                            "buffered); total norm read back on the host every step"},
            "gpu_launches": launches,
            "nccl_allreduces": 0 if ddp._fused else allreduces,
            "fused_allreduces": stats1.get("fused_allreduces", 0) - stats0.get("fused_allreduces", 0),
            "allreduce_impl": ("fused peer-memory kernel (" + ("multicast" if any(a.use_multicast for a in ddp._sym.values()) else "peer") + ")")
                              if ddp._fused else ((fused_note or "nccl") if ws > 1 else "none"),
            "roofline": {"kernel": ("sharded_adamw_kernel (AdamW on the owned chunks + parameter all-gather, bf16)" if zero1 else
                                    "adamw_bulk_kernel (fused clip+AdamW, bf16, TMA-staged inputs)"), "bound": "hbm", "achieved": achieved,
                         "peak": peak, "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src,
                         "frac_of_8000_nominal": achieved / 8000.0, "algorithmic_bytes_per_launch": alg_bytes,
                         "avg_launch_ms": adamw_ms, "traffic": adamw_traffic_from_profile()},
            "path_bytes_per_step": {"pack": 0 if ws == 1 else 4 * total, "norm": 2 * total, "adamw": 14 * total},
            "path_gbs": ((0 if ws == 1 else 4 * total) + 2 * total + 14 * total) / (ms / args.steps * 1e-3) / 1e9,
            "clocks": clk, "total_norm": last_norm,
            "gradients": ("resident in the bucket arena (p.grad is its bucket view): no pack launch, the 1/world_size "
                          "scale is applied inside the collective") if in_arena and ddp._fused and stats1.get("folded_scale_launches", 0)
                         else ("separate buffer: one pack (copy + scale) per step" if ws > 1 else "in place, nothing exchanged"),
        }
        if parity is not None:
            out["parity_check"] = parity
        if ws > 1:
            out["per_bucket_path"] = {"ms_per_step": per_bucket_ms, "steps": 5, "buckets": len(bucket_bytes),
                                      "note": "the same step with one collective per bucket, launched in bucket order as the "
                                              "autograd hooks of a real backward do; not hidden behind a backward here"}
        if ws > 1:
            # bucket allreduce bus bandwidth, measured alone on the same buckets (NCCL-tests convention)
            out["allreduce"] = measure_allreduce(ddp, ws, dev)
    elif ws > 1:
        measure_allreduce(ddp, ws, dev)

    if rank == 0 and ws == 1 and not args.no_side_records and args.workload == "llama-3.2-1b":
        del ddp, opt
        torch.cuda.empty_cache()
        try:
            out["incumbent"] = incumbent_record(params, grads, dev)
        except Exception as e:
            out["incumbent"] = {"error": f"{type(e).__name__}: {str(e)[:200]}"}
        del model, params, grads, gflat
        torch.cuda.empty_cache()
        try:
            out["train"] = train_record(dev)
        except Exception as e:
            out["train"] = {"error": f"{type(e).__name__}: {str(e)[:200]}"}
    if rank == 0 and ws == 1 and not args.no_cpu_baseline:
        os.sched_setaffinity(0, all_cpus)            # the CPU baseline gets every host core
        out["cpu_baseline"] = cpu_baseline_sample(cfg, total)
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.destroy_process_group()
    if parity is not None and not parity["ok"]:
        sys.exit(3)                                  # the launcher's exit code carries a failed self-check


def measure_allreduce(ddp, ws, dev):
    """Bus bandwidth of the path's gradient collective, measured alone: busbw = S/t * 2(n-1)/n
    (NCCL-tests convention). `coalesced` is what a step of this bench issues (all buckets complete at
    once -> one allreduce over the contiguous bucket arena); `per_bucket` is what a real backward
    issues (one allreduce per 25 MiB-capped bucket, back to back on one stream)."""
    def timed(fn, reps=5):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b) / reps], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    bufs = [bk.buffer for bk in ddp._buckets]
    arenas = list(ddp._arenas.values())
    S = sum(b.numel() * b.element_size() for b in bufs)
    Sa = sum(a.numel() * a.element_size() for a in arenas)
    ms_one = timed(lambda: [dist.all_reduce(a) for a in arenas])
    ms_per = timed(lambda: [dist.all_reduce(b) for b in bufs])
    fused = None
    if ddp._fused:
        # the path's own collective: allreduce AND per-tensor sums of squares in one kernel (+2 barriers + fold)
        import ddp_b200
        k = ddp._kernels
        ss = torch.empty_like(ddp._sumsq)
        tab = ddp._view_table

        def run_fused():
            for bk_dt, arena in ddp._sym.items():
                rows = [bk for bk in ddp._buckets if bk.dtype == bk_dt]
                first, n = rows[0].first, sum(bk.n for bk in rows)
                k.allreduce_sumsq(tab, bk_dt, arena, ss[first:first + n], first=first, count=n, scatter=False)
        if len(ddp._sym) == 1:
            ms_f = timed(run_fused)
            mc = any(a.use_multicast for a in ddp._sym.values())
            # bytes one GPU really moves per NVLink direction: multicast (1 + 1/n) S (its copy is read once by
            # the switch + it stores its 1/n through the switch), peer 2(n-1)/n S (as a ring)
            wire = Sa * ((1 + 1 / ws) if mc else 2 * (ws - 1) / ws)
            fused = {"ms": ms_f, "busbw_equiv_gbs": Sa / (ms_f * 1e-3) * 2 * (ws - 1) / ws / 1e9,
                     "busbw_convention": "NCCL-tests ring equivalent, S/t * 2(n-1)/n",
                     "wire_gbs_per_direction": wire / (ms_f * 1e-3) / 1e9,
                     "wire_bytes_per_direction": wire, "wire_frac_of_900_nominal": wire / (ms_f * 1e-3) / 1e9 / 900.0,
                     "wire_frac_of_770_measured_peer_copy": wire / (ms_f * 1e-3) / 1e9 / 770.0,
                     "mode": "multicast" if mc else "peer",
                     "includes": "allreduce + per-tensor sum of squares + 2 barriers + fold"}
    for a in arenas:
        a.zero_()
    f = 2 * (ws - 1) / ws / 1e9
    return {"bytes_per_step": S, "buckets": len(bufs), "fused_kernel": fused,
            "coalesced": {"ms": ms_one, "busbw_gbs": Sa / (ms_one * 1e-3) * f, "frac_of_900": Sa / (ms_one * 1e-3) * f / 900.0},
            "per_bucket": {"ms": ms_per, "busbw_gbs": S / (ms_per * 1e-3) * f, "frac_of_900": S / (ms_per * 1e-3) * f / 900.0}}


# ---- CPU baselines --------------------------------------------------------------------------------
def _stock_cpu_worker(r, n, store, cfg, steps, warmup, threads, budget_s, out_path):
    """One rank of the reference's own CPU path: UNMODIFIED stock torch — DistributedDataParallel on gloo
    + clip_grad_norm_ + torch.optim.AdamW — on the FULL parameter set, gradients delivered through
    autograd (so the Reducer's hooks, bucket copies and the gloo allreduce all run as in training)."""
    torch.set_num_threads(threads)
    dist.init_process_group("gloo", init_method=f"file://{store}", rank=r, world_size=n)
    shapes = llama_param_shapes(**cfg)
    model = ParamSet(shapes, torch.bfloat16, torch.device("cpu"))
    params = list(model.parameters())
    total = sum(p.numel() for p in params)
    ddp = torch.nn.parallel.DistributedDataParallel(model)
    opt = torch.optim.AdamW(params, lr=LR, weight_decay=1e-2)
    gen = torch.Generator().manual_seed(1 + r)
    holder = {"grads": [torch.empty_like(p).normal_(0, 0.01, generator=gen) for p in params]}
    # ParamSet has no forward of its own: the synthetic-gradient function stands in for the model
    model.forward = lambda: SynthGrad.apply(holder, *params)

    def one():
        loss = ddp()
        loss.backward()
        torch.nn.utils.clip_grad_norm_(params, MAX_NORM)
        opt.step()
        opt.zero_grad()

    t0 = time.perf_counter()
    warmed = 0
    for _ in range(warmup):
        one()
        warmed += 1
        # bounded runs (many ranks on few cores): stop warming up once it has cost 20 s; every rank decides alike
        slow = torch.tensor([1.0 if (budget_s is not None and time.perf_counter() - t0 > 20.0) else 0.0])
        dist.all_reduce(slow, op=dist.ReduceOp.MAX)
        if float(slow) > 0:
            break
    per = torch.tensor([(time.perf_counter() - t0) / max(1, warmed)], dtype=torch.float64)
    dist.all_reduce(per, op=dist.ReduceOp.MAX)
    # the whole run has to end within minutes: with more ranks than the host has cores to give them, a step
    # takes many seconds, and the step count is capped (every rank computes the same cap)
    run = steps if budget_s is None else max(2, min(steps, int(budget_s / max(float(per), 1e-3))))
    dist.barrier()
    t0 = time.perf_counter()
    for _ in range(run):
        one()
    dist.barrier()
    dt = time.perf_counter() - t0
    if r == 0:
        json.dump({"rate": run / dt, "steps": run, "total": total, "threads": threads, "warmup_s_per_step": float(per),
                   "warmup": warmed},
                  open(out_path, "w"))
    dist.destroy_process_group()


def stock_cpu_step_rate(cfg, steps, warmup, ranks=1, budget_s=None):
    """-> dict(rate = steps/s of ONE rank's loop, steps actually timed, total elements, threads per rank).
    ranks > 1: this process spawns a gloo group of that many workers on the host cores."""
    import tempfile
    import torch.multiprocessing as mp
    d = tempfile.mkdtemp(prefix="ddpb200_ref_")
    store, out_path = os.path.join(d, "store"), os.path.join(d, "out.json")
    cores = max(1, len(os.sched_getaffinity(0)))
    threads = max(1, cores // ranks)
    # a file store of its own: under torchrun both env:// and tcp:// would try to join the elastic agent's store
    if ranks == 1:
        assert not dist.is_initialized()
        _stock_cpu_worker(0, 1, store, cfg, steps, warmup, threads, budget_s, out_path)
    else:
        mp.spawn(_stock_cpu_worker, args=(ranks, store, cfg, steps, warmup, threads, budget_s, out_path), nprocs=ranks, join=True)
    return json.load(open(out_path))


def host_mem_available_gb():
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable:"):
                return int(line.split()[1]) / 1e6
    except OSError:
        pass
    return None


def reference_workload_string(workload, n_tensors, total):
    return f"{workload}-arch parameter set, bf16 p/g/m/v, {n_tensors} tensors, {total} elements"


def workload_config(workload, n_tensors, total, ws):
    """The `config` of the JSON line: what is computed, the same dictionary for our arm and for the reference
    arm (how each arm computes it is said outside `config`)."""
    return {"workload": reference_workload_string(workload, n_tensors, total),
            "path": f"DDP gradient sync (25 MiB buckets, averaged over the replicas) + clip_grad_norm_(max_norm {MAX_NORM}) "
                    f"+ AdamW(lr {LR}, weight_decay 0.01), one step per synthetic gradient set",
            "global_batch": f"{ws} replicas x 1 synthetic gradient set",
            "l2": f"working set per step {16 * total / 1e9:.1f} GB >> 126 MB L2; no flush needed",
            "parallelism": f"dp{ws}"}


def cpu_baseline_sample(cfg, total_full):
    """Bounded sample beside our N=1 number: the stock torch CPU path on the FULL parameter set, 5 steps
    after 1 warm-up, every host core. (Runs in a child process: this one already holds an NCCL group.)"""
    r = stock_cpu_step_rate_subprocess(cfg, steps=5, warmup=1)
    return {"value": r["rate"], "unit": UNIT, "cores": r["threads"], "kind": "reference",
            "sample": f"stock torch DDP(gloo,1 rank)+clip_grad_norm_+AdamW on CPU, bf16, the full parameter set "
                      f"({r['total']} elements), {r['steps']} steps after 1 warm-up",
            "host_cpus": os.cpu_count()}


def stock_cpu_step_rate_subprocess(cfg, steps, warmup):
    """stock_cpu_step_rate(ranks=1) in a fresh interpreter with every host core (this process may be bound
    to the GPU's NUMA node and carries torchrun's OMP_NUM_THREADS=1)."""
    code = ("import json,sys,os; sys.path.insert(0, %r); import torch, bench; "
            "torch.set_num_threads(max(1, len(os.sched_getaffinity(0)))); "
            "print('RESULT ' + json.dumps(bench.stock_cpu_step_rate(json.loads(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]))))" % ROOT)
    env = {k: v for k, v in os.environ.items() if not (k.startswith("TORCHELASTIC") or k in (
        "OMP_NUM_THREADS", "RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT", "GROUP_RANK", "LOCAL_WORLD_SIZE"))}
    env["CUDA_VISIBLE_DEVICES"] = ""
    res = subprocess.run([sys.executable, "-c", code, json.dumps(cfg), str(steps), str(warmup)], env=env,
                         stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=900)
    for line in res.stdout.splitlines():
        if line.startswith("RESULT "):
            return json.loads(line[7:])
    raise RuntimeError("CPU baseline child failed: " + res.stderr[-400:])


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on the host cores, on our arm's
    workload (the full parameter set), honouring --steps/--warmup. Under torchrun rank 0 alone works (the
    other ranks exit 0); at N>1 it drives an N-rank gloo group of its own — a DDP step's CPU path has an
    allreduce — when the host has the memory for N replicas (about 13 GB each), else one rank, and says which."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    cfg = WORKLOADS[args.workload]
    shapes = llama_param_shapes(**cfg)
    total_full = sum(int(torch.Size(s).numel()) for s in shapes)
    # every host core: torchrun exports OMP_NUM_THREADS=1 to its workers, which is not the reference's best
    os.environ.pop("OMP_NUM_THREADS", None)
    torch.set_num_threads(max(1, len(os.sched_getaffinity(0))))
    ranks = max(1, args.gpus)
    need_gb = ranks * 10.5 * total_full * 1e-9 + 8        # p + g + bucket + two states in bf16, plus slack
    mem = host_mem_available_gb()
    note = None
    if ranks > 1 and mem is not None and mem < need_gb:
        note = f"host has {mem:.0f} GB available, {ranks} CPU replicas need ~{need_gb:.0f} GB: 1-rank group instead"
        ranks = 1
    for k in [k for k in os.environ if k.startswith("TORCHELASTIC")]:
        os.environ.pop(k)
    r = stock_cpu_step_rate(cfg, steps=max(1, args.steps), warmup=max(1, args.warmup), ranks=ranks,
                            budget_s=None if ranks == 1 else 175.0)
    value = r["rate"] * ranks
    sample = (f"stock torch DDP(gloo, {ranks} rank{'s' if ranks > 1 else ''})+clip_grad_norm_+AdamW(default) on CPU, bf16, "
              f"the full parameter set ({r['total']} elements per rank), {r['steps']} steps after {r['warmup']} warm-up, "
              f"{r['threads']} threads per rank")
    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": r["steps"], "warmup": r["warmup"],
           "ms_per_step": 1e3 / r["rate"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "bf16", "data": "synthetic", "impl": "reference",
           "config": workload_config(args.workload, len(shapes), total_full, max(1, args.gpus)),
           "implementation": f"stock torch on the host cores: DistributedDataParallel (gloo, {ranks} rank{'s' if ranks > 1 else ''}) "
                             f"+ clip_grad_norm_ + torch.optim.AdamW",
           "steps_cap": None if r["steps"] == max(1, args.steps) else
                        f"{r['steps']} of the {args.steps} requested steps timed (a step takes {r['warmup_s_per_step']:.1f} s "
                        f"with {r['threads']} threads per rank; run bounded to a few minutes)",
           "note": note,
           "cpu_baseline": {"value": value, "unit": UNIT, "cores": r["threads"] * ranks, "kind": "reference", "sample": sample,
                            "host_cpus": os.cpu_count()},
           "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=list(WORKLOADS), default="llama-3.2-1b")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-side-records", action="store_true",
                    help="N=1: skip the `train` (whole training step vs stock torch) and `incumbent` (torch's CUDA "
                         "kernels on the same tensors) records")
    ap.add_argument("--no-numa-bind", action="store_true", help="do not bind the rank to its GPU's NUMA-local CPUs")
    ap.add_argument("--skip-e2e", action="store_true", help="skip the host-gradient leg (large workloads: pinned host memory)")
    ap.add_argument("--separate-grads", action="store_true",
                    help="N>1: keep the synthetic gradients in a buffer of their own (one pack launch per step) instead "
                         "of in the bucket arena")
    ap.add_argument("--no-parity-check", action="store_true", help="N>1: skip the self-check before the timed loop")
    ap.add_argument("--no-zero1-check", action="store_true", help="N>1: skip part (iv) of the self-check (sharded update)")
    ap.add_argument("--zero1", action="store_true",
                    help="N>1: the sharded update (reduce-scatter + AdamW on owned chunks + all-gather of parameters) "
                         "instead of the drop-in replicated path; an extension, not the default")
    ap.add_argument("--allreduce", choices=["nccl", "fused"], default=os.environ.get("DDPB200_BENCH_ALLREDUCE", "fused"),
                    help="N>1: NCCL allreduce + norm launch, or the fused peer-memory allreduce+norm kernel")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
