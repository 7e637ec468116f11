#!/usr/bin/env python
"""Throughput bench for the B200-native LightGlue matcher forward.

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU path (oracle port)

Metric (BASELINE.json): image pairs/sec at 2048 keypoints/image.  One step = one forward over
a batch of 32 synthetic pairs x 2048 SuperPoint-shaped keypoints (BASELINE config 2) per GPU;
pairs shard across ranks with no data-path collective (weak scaling).  Prints ONE JSON line.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

if "reference" in sys.argv[1:]:
    # torchrun exports OMP_NUM_THREADS=1 to every rank; the CPU arm is meant to use all host cores
    for _k in ("OMP_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[_k] = str(os.cpu_count() or 1)

# stdout carries exactly ONE JSON line: everything else a library prints there (NCCL's version banner under
# NCCL_DEBUG=VERSION, ...) is sent to stderr by pointing fd 1 at fd 2 until the line is written
_JSON_FD = os.dup(1)
os.dup2(2, 1)

import torch


def emit(line: dict):
    os.write(_JSON_FD, (json.dumps(line) + "\n").encode())

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "cs566-project-lightglue_b200"
METRIC, UNIT = "image pairs/sec @2048 kp/img", "pairs/s"
CONF = {"filter_threshold": 0.1}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "eager"])
    ap.add_argument("--batch", type=int, default=32, help="pairs per GPU per step")
    ap.add_argument("--kpts", type=int, default=2048)
    ap.add_argument("--cpu-pairs", type=int, default=4, help="pairs timed for the cpu_baseline sample")
    ap.add_argument("--flash", action="store_true",
                    help="--impl eager: the reference's flash=True branch (fp16 SDPA per direction, TF32 allowed)")
    ap.add_argument("--no-eager", action="store_true", help="skip the gpu_eager_baseline legs of the b200 arm")
    return ap.parse_args()


def workload_name(a):
    return f"SuperPoint+LightGlue (9 layers, d=256, 4 heads), {a.batch} pairs x {a.kpts} keypoints per GPU, synthetic"


def config_dict(a, world):
    """The SAME config object in every arm (the driver compares them key by key)."""
    return {"workload": workload_name(a), "pairs_per_gpu": a.batch, "keypoints": a.kpts,
            "l2": "inputs+activations per step (>4 GB) exceed the 126 MB L2; no flush needed",
            "parallelism": f"pairs sharded over {world} GPU(s), no data-path collective"}


# ----------------------------------------------------------------------------- CPU arm (oracle port)
def cpu_forward_rate(kpts: int, pairs: int, warm: int = 1):
    """pairs/s of the reference algorithm on the host cores: the oracle port (torch CPU ops in the
    reference's op order, bit-identical to the reference module in the build container), B=1 per
    forward as the reference's eval loop does (utils/export_predictions.py:32)."""
    from oracle import lightglue_oracle as O

    W = importlib.import_module(PKG + ".weights")
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sd = W.make_state_dict(CONF, seed=0)
    data = W.make_inputs(1, kpts, kpts, seed=1)
    for _ in range(warm):
        O.forward(sd, CONF, data)
    t0 = time.perf_counter()
    for _ in range(pairs):
        O.forward(sd, CONF, data)
    dt = time.perf_counter() - t0
    return pairs / dt, cores, dt


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # each step = a bounded sample of the workload: ONE pair of the 32-pair batch
    for _ in range(max(a.warmup, 1)):
        rate, cores, _ = cpu_forward_rate(a.kpts, 1, warm=0)
    rate, cores, dt = cpu_forward_rate(a.kpts, a.steps, warm=0)
    sample = f"{a.steps} forwards of 1 pair x {a.kpts} kpts (1/{a.batch} of the step's batch), fp32, {cores} threads"
    emit({
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": 1e3 * dt / a.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(a, a.gpus),
        "note": "CPU arm: oracle port of the reference (the Python reference cannot travel to the GPU box); a step is a "
                "bounded sample of the workload: one pair of the batch",
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    })


def _flash_oracle(O):
    """The reference's flash=True branch on the oracle's op sequence: Attention.forward casts q, k, v to fp16 and calls
    SDPA (lightglue.py:139-145); CrossBlock runs one such call per direction (:229-233) instead of the shared
    similarity.  Returns a restore function."""
    import torch.nn.functional as F

    keep = (O.attention, O.cross_block)

    def attention(q, k, v):
        return F.scaled_dot_product_attention(q.half().contiguous(), k.half().contiguous(), v.half().contiguous()).to(q.dtype)

    def cross_block(sd, p, x0, x1, heads=4):
        def split(t):
            return t.unflatten(-1, (heads, -1)).transpose(1, 2)

        qk0, qk1 = split(O._lin(sd, p + ".to_qk", x0)), split(O._lin(sd, p + ".to_qk", x1))
        v0, v1 = split(O._lin(sd, p + ".to_v", x0)), split(O._lin(sd, p + ".to_v", x1))
        m0, m1 = attention(qk0, qk1, v1), attention(qk1, qk0, v0)
        m0 = O._lin(sd, p + ".to_out", m0.transpose(1, 2).flatten(start_dim=-2))
        m1 = O._lin(sd, p + ".to_out", m1.transpose(1, 2).flatten(start_dim=-2))
        return O._ffn(sd, p, x0, m0), O._ffn(sd, p, x1, m1)

    O.attention, O.cross_block = attention, cross_block

    def restore():
        O.attention, O.cross_block = keep

    return restore


def eager_rate(batch, kpts, steps, warmup, flash, dev):
    """pairs/s of the reference's op sequence in PyTorch eager on the SAME GPU: the oracle port on cuda tensors.
    flash=False: fp32 SDPA as in lightglue.py:146-149, TF32 off (the reference's most accurate GPU branch);
    flash=True: its fastest one -- fp16 SDPA per direction (:139-145, :229-233), TF32 matmuls allowed."""
    from oracle import lightglue_oracle as O

    W = importlib.import_module(PKG + ".weights")
    tf32 = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
    torch.backends.cuda.matmul.allow_tf32 = torch.backends.cudnn.allow_tf32 = bool(flash)
    restore = _flash_oracle(O) if flash else (lambda: None)
    try:
        sd = {k: v.to(dev) for k, v in W.make_state_dict(CONF, seed=0).items()}
        data = to_device(W.make_inputs(batch, kpts, kpts, seed=1), dev, non_blocking=False)
        with torch.no_grad():
            for _ in range(warmup):
                O.forward(sd, CONF, data)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                O.forward(sd, CONF, data)
            e1.record()
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    finally:
        restore()
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = tf32
    return batch * steps / (ms / 1e3), ms / steps


def run_eager(a):
    """Informative arm (SURVEY 8(d) "extra informative baseline"): the reference's algorithm in PyTorch eager on the
    SAME B200.  Not the product and not the reference arm the driver compares against."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    dev = torch.device("cuda", 0)
    rate, ms = eager_rate(a.batch, a.kpts, a.steps, max(a.warmup, 3), a.flash, dev)
    emit({
        "impl": "eager", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": 1,
        "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None,
        "dtype": "f16 SDPA per direction, TF32 matmuls (PyTorch eager, the reference's flash=True branch)" if a.flash
                 else "f32 (PyTorch eager, cuBLAS / SDPA, TF32 off)",
        "data": "synthetic", "config": config_dict(a, 1),
        "note": "oracle port of the reference on cuda tensors: informative only", "gpu_launches": 0,
    })


# ----------------------------------------------------------------------------- clocks sampler
class ClockSampler(threading.Thread):
    """SM clock + throttle reasons DURING the timed region: NVML in-process (a sample every 5 ms), nvidia-smi as the
    fallback when pynvml is missing (one sample per ~100 ms: short runs on busy boxes may then stay unsampled)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    BITS = {"sw_power_cap": 0x4, "hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], False
        self.nvml = self.handle = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nvml, self.handle = pynvml, pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nvml = None

    def _reasons(self):
        for fn in ("nvmlDeviceGetCurrentClocksEventReasons", "nvmlDeviceGetCurrentClocksThrottleReasons"):
            if hasattr(self.nvml, fn):
                return int(getattr(self.nvml, fn)(self.handle))
        return 0

    def run(self):
        while not self.stop_flag:
            try:
                if self.nvml is not None:
                    mhz = self.nvml.nvmlDeviceGetClockInfo(self.handle, self.nvml.NVML_CLOCK_SM)
                    bits = self._reasons()
                    flags = ["Active" if bits & self.BITS[n] else "Not Active" for n in
                             ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")]
                    self.samples.append([str(mhz), str(self.max_mhz)] + flags)
                    time.sleep(0.005)
                    continue
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                if len(f) >= 6:
                    self.samples.append(f)
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        mhz = sorted(int(s[0]) for s in self.samples if s[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(s[2 + i].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None,
                "sm_max_mhz": int(self.samples[0][1]) if self.samples[0][1].isdigit() else None,
                "reasons": reasons, "samples": len(self.samples),
                "source": "nvml" if self.nvml is not None else "nvidia-smi"}


# ----------------------------------------------------------------------------- GPU arm
def to_device(data, dev, non_blocking=True):
    return {k: ({kk: vv.to(dev, non_blocking=non_blocking) for kk, vv in v.items()} if isinstance(v, dict)
                else v.to(dev, non_blocking=non_blocking)) for k, v in data.items()}


def pin(data):
    return {k: ({kk: vv.pin_memory() for kk, vv in v.items()} if isinstance(v, dict) else v.pin_memory())
            for k, v in data.items()}


def nbytes(data):
    return sum((sum(vv.numel() * vv.element_size() for vv in v.values()) if isinstance(v, dict)
                else v.numel() * v.element_size()) for v in data.values())


def run_b200(a):
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py (impl b200) needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    distributed = world > 1
    if distributed:
        dist.init_process_group("nccl", device_id=dev)
    C = importlib.import_module(PKG + "._cabi")
    M = importlib.import_module(PKG + ".matcher")
    W = importlib.import_module(PKG + ".weights")
    P = importlib.import_module(PKG + ".pipeline")
    model = M.LightGlue(CONF)
    model.load_state_dict(W.make_state_dict(CONF, seed=0), strict=True)
    model = model.to(dev)
    host = pin(W.make_inputs(a.batch, a.kpts, a.kpts, seed=1 + rank))  # each rank owns its own pairs
    data = to_device(host, dev)
    torch.cuda.synchronize()

    def barrier():
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if distributed:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(max(a.warmup, 3)):
        model(data)

    # ---- end to end through the public API with HOST buffers ("e2e")
    out_host = {k: torch.empty((a.batch, a.kpts), dtype=dt).pin_memory()
                for k, dt in (("matches0", torch.int64), ("matches1", torch.int64),
                              ("matching_scores0", torch.float32), ("matching_scores1", torch.float32))}

    def step_e2e():
        pred = model(to_device(host, dev))
        for k, buf in out_host.items():
            buf.copy_(pred[k], non_blocking=True)
        torch.cuda.current_stream().synchronize()  # the caller reads the matches (export_predictions.py:68)

    for _ in range(2):
        step_e2e()
    ms_e2e_sync = timed(step_e2e, a.steps)

    # the same loop through HostPipeline: copies of batch i+1 / i-1 on their own streams under forward(i);
    # every batch is still copied in from pinned host memory and its matches copied back inside the timed region
    pipe = P.HostPipeline(model, depth=2)

    def stream_e2e(n):
        got = 0
        for pred in pipe.run(host for _ in range(n)):
            got += int(pred["matches0"].shape[0])
        assert got == n * a.batch

    stream_e2e(max(a.warmup, 3) + 17)  # also lets a fresh box ramp its clocks before the first timed leg
    sampler = ClockSampler(local)  # clocks / throttle reasons over both timed legs (e2e, value)
    sampler.start()
    ms_e2e = timed(lambda: stream_e2e(a.steps), 1)
    e2e = world * a.batch * a.steps / (ms_e2e / 1e3)

    # ---- device-resident throughput ("value")
    def step():
        return model(data)

    for _ in range(max(a.warmup, 3)):
        step()
    l0 = C.LAUNCHES
    ms = timed(step, a.steps)
    launches = C.LAUNCHES - l0
    sampler.stop_flag = True
    value = world * a.batch * a.steps / (ms / 1e3)

    # ---- self-check outside the timed regions: three pairs of the timed batch (several work items per persistent CTA)
    # against the same pairs one per call (one item per CTA); a synchronisation fault at scale shows here as ~5e-2
    pred = step()
    la_batch = pred["log_assignment"]
    check_err, check_same = 0.0, True
    for b in sorted({0, a.batch // 2, a.batch - 1}):
        one = model({k: ({kk: vv[b:b + 1] for kk, vv in v.items()} if isinstance(v, dict) else v[b:b + 1])
                     for k, v in data.items()})
        check_err = max(check_err, float((one["log_assignment"][0] - la_batch[b]).abs().max()))
        check_same = check_same and bool(torch.equal(one["matches0"][0], pred["matches0"][b]))
    # and the e2e path: what HostPipeline hands back for the same host batch (copies on side streams, graph replays)
    pipe_same = True
    for got in pipe.run(host for _ in range(3)):
        pipe_same = pipe_same and bool(torch.equal(got["matches0"], pred["matches0"].cpu())) and \
            bool(torch.equal(got["matching_scores0"], pred["matching_scores0"].cpu()))
    check = {"max_abs_diff_log_assignment_vs_one_pair_calls": check_err, "matches0_identical": check_same,
             "host_pipeline_identical": pipe_same}
    if not pipe_same:
        raise RuntimeError("bench self-check failed: HostPipeline results differ from the direct call")
    if check_err > 1e-3:
        raise RuntimeError(f"bench self-check failed: batched forward differs from one-pair calls by {check_err}")

    # ---- roofline of the dominant kernel (attention): CUDA events around every launch
    C.ATTN_EVENTS = []
    barrier()
    for _ in range(min(a.steps, 3)):
        step()
    torch.cuda.synchronize()
    ev = C.ATTN_EVENTS
    C.ATTN_EVENTS = None
    attn_ms = sum(e0.elapsed_time(e1) for e0, e1, _ in ev)
    attn_flops = sum(f for _, _, f in ev)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("bf16_tflops_sustained", 1400.0))  # kernel timed inside a long step
    peak_burst = float(peaks.get("bf16_tflops", 1650.0))
    achieved = attn_flops / (attn_ms / 1e3) / 1e12
    traffic = None  # DRAM bytes per launch of this kernel from the committed ncu --set full capture
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "r2_attention_traffic.json")))["dram_bytes_per_launch"]
    except Exception:
        pass
    roofline = {"kernel": "attention_kernel (tcgen05 flash attention, self + cross)", "bound": "tensor",
                "flops": "algorithmic: 4*64*nq*nk per self launch, 3*64*nq*nk per cross direction (S counted once)",
                "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak, "frac_of_burst_peak": achieved / peak_burst,
                "traffic": traffic,
                "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained" if peaks else "fallback 1.4 PFLOP/s sustained",
                "launches_timed": len(ev), "avg_launch_ms": attn_ms / max(len(ev), 1),
                "share_of_step": attn_ms / max(min(a.steps, 3), 1) / (ms / a.steps)}

    # ---- the HBM-bound pass: assignment + filter on the engine left resident by the last step (final_proj tiles,
    # matchability logits, strip pass 1, reduce, strip pass 2 with keys, filter).  Algorithmic bytes per pair
    # (SURVEY.md 8(d)): write (M+1)(N+1)*4 + read 2*N*256*4 of projected descriptors.
    eng = model._engine(a.batch, a.kpts, a.kpts, dev, cached=True)
    rows0, rows1 = eng.rows()
    n_layers = model.conf.n_layers

    def assign_step():
        model._assign(n_layers - 1, eng, rows0, rows1, None, None, th=model.conf.filter_threshold)

    for _ in range(3):
        assign_step()
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(10):
        assign_step()
    ev1.record()
    torch.cuda.synchronize()
    assign_ms = ev0.elapsed_time(ev1) / 10
    assign_bytes = a.batch * ((a.kpts + 1) ** 2 * 4 + 2 * a.kpts * 256 * 4)
    hbm_peak = float(peaks.get("hbm_gbs", 6500.0))
    assign_gbs = assign_bytes / (assign_ms / 1e3) / 1e9
    roofline_assign = {"kernel": "assign_f16_kernel<1> (rows) + <1> (columns) + merge + assign_f16_kernel<2> + filter_finalize (final_proj and logits included)",
                       "bound": "hbm", "achieved": assign_gbs, "peak": hbm_peak, "unit": "GB/s",
                       "frac": assign_gbs / hbm_peak, "ms": assign_ms, "algorithmic_bytes": assign_bytes,
                       "share_of_step": assign_ms / (ms / a.steps)}

    # ---- the reference's op sequence in PyTorch eager on this same GPU (informative; N=1 only): its most accurate
    # branch (fp32, TF32 off) and its fastest one (flash=True: fp16 SDPA, TF32 on), at the bench batch and one pair per call
    eager = None
    if rank == 0 and world == 1 and not a.no_eager:
        del eng, rows0, rows1, pred, la_batch
        model.__dict__.pop("_engines", None)
        torch.cuda.empty_cache()
        eager = {"unit": UNIT, "what": "oracle port of the reference on cuda tensors, PyTorch eager (cuBLAS / SDPA)"}
        for name, flash in (("fp32_tf32_off", False), ("flash_fp16_sdpa_tf32_on", True)):
            r32, ms32 = eager_rate(a.batch, a.kpts, 3, 2, flash, dev)
            r1, ms1 = eager_rate(1, a.kpts, 10, 3, flash, dev)
            eager[name] = {"pairs_per_s_batch": r32, "ms_per_step_batch": ms32, "pairs_per_s_one_pair_per_call": r1,
                           "ms_one_pair": ms1}
        torch.cuda.empty_cache()

    if rank == 0:
        cpu = None  # timed at N=1 only (under torchrun the other ranks' host threads would share the cores)
        if world == 1:
            cpu_rate, cores, cpu_dt = cpu_forward_rate(a.kpts, a.cpu_pairs)
            cpu = {"value": cpu_rate, "unit": UNIT, "cores": cores, "kind": "port",
                   "sample": f"{a.cpu_pairs} forwards of 1 pair x {a.kpts} kpts ({cpu_dt:.1f} s), fp32 torch CPU"}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
            "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16 attention operands, 3-term split-bf16 linears, fp16 similarity operands, f32 accumulate / softmax / LayerNorm", "data": "synthetic",
            "config": config_dict(a, world),
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": nbytes(host),
                    "d2h_bytes_per_step": sum(v.numel() * v.element_size() for v in out_host.values()),
                    "ms_per_step": ms_e2e / a.steps, "api": "HostPipeline(model).run(host batches)",
                    "unpipelined_ms_per_step": ms_e2e_sync / a.steps},
            "gpu_launches": launches,
            "roofline": roofline,
            "roofline_assignment": roofline_assign,
            "cpu_baseline": cpu,
            "gpu_eager_baseline": eager,
            "check": check,
            "clocks": sampler.summary(),
        }
        emit(line)
    if distributed:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    elif args.impl == "eager":
        run_eager(args)
    else:
        run_b200(args)
