#!/usr/bin/env python
"""bench.py -- RVQ encode & decode throughput (BASELINE.json metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A step = one pass of the hot path over one batch of synthetic latents: `encode(x)` followed by
`decode(codes)` through the drop-in module (audiocraft_b200.quantization.ResidualVectorQuantizer),
i.e. through the C ABI.  At N = 1 the workload is BASELINE.json configs[1] (MusicGen 32 kHz RVQ,
n_q=4 bins=2048 dim=128, B=64 x T=1500).  For N > 1 (torchrun, one rank per GPU) every rank
encodes+decodes its own batch of that shape (frames are independent: weak scaling, no collective
in the data path) and the codes of every step are exchanged between the ranks (peer stores into
symmetric memory over NVLink, or one NCCL all-gather) inside the timed region, as
BASELINE.json configs[4] describes.

Rank 0 prints ONE JSON line.  `value` is frames/s of the whole job with inputs resident in HBM;
`e2e` is the same through the public API with pinned HOST buffers (H2D of the latents, D2H of
codes and quantized latents inside the timed region); `roofline` is the encode kernel against the
measured tensor peak; `cpu_baseline` is the staged reference on the box's host cores (the C port beside it).

`--impl reference` times the reference's own CPU implementation with every host thread: the unmodified
`audiocraft.quantization` files staged in the git-ignored `oracle/_ref/` by `oracle/fetch_ref.py` (kind
"reference", the full batch per step); only when that copy is absent, the C port of `oracle/` (kind "port").
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "RVQ encode & decode Mframes/s"
UNIT = "Mframes/s"
CFG = dict(D=128, C=2048, K=4, B=64, T=1500)  # BASELINE.json configs[1]
WORKLOAD = "configs[1]: MusicGen 32 kHz EnCodec RVQ n_q=4 bins=2048 dim=128, B=64 x T=1500 frames, encode+decode"
N_ROTATING = 4  # 4 x 49 MB inputs (197 MB) > 126 MB L2


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# Rank 0 must print exactly one JSON line on stdout.  Libraries (NCCL's version banner, for one) also write
# to file descriptor 1, so the real stdout is set aside and fd 1 is pointed at stderr for the rest of the run.
_REAL_STDOUT = None


def claim_stdout():
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line: dict):
    out = _REAL_STDOUT if _REAL_STDOUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return dict(bf16=float(p["bf16_tflops"]), bf16_sustained=float(p.get("bf16_tflops_sustained", 0) or 0),
                    hbm=float(p["hbm_gbs"]), source="measured (MEASURED_PEAKS.json)")
    return dict(bf16=1590.0, bf16_sustained=1400.0, hbm=6650.0, source="fallback (B200_PROFILING.md)")


# ------------------------------------------------------------------------------------------------
# clock sampling during the timed region (NVML; nvidia-smi as a fallback)
# ------------------------------------------------------------------------------------------------
_SAMPLER_SRC = r"""
import sys, time
import pynvml as nv
nv.nvmlInit()
bus = sys.argv[1]
h = None
for i in range(nv.nvmlDeviceGetCount()):
    hh = nv.nvmlDeviceGetHandleByIndex(i)
    b = nv.nvmlDeviceGetPciInfo(hh).busId
    b = b.decode() if isinstance(b, bytes) else b
    if bus and b.lower().endswith(bus.lower()):
        h = hh
if h is None:
    h = nv.nvmlDeviceGetHandleByIndex(int(sys.argv[2]))
print("max", nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM), flush=True)
while True:
    try:
        print(repr(time.time()), nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM),
              nv.nvmlDeviceGetCurrentClocksThrottleReasons(h), flush=True)
    except Exception:
        pass
    time.sleep(0.002)
"""


class ClockSampler:
    """SM clock and throttle reasons every 2 ms, from a SEPARATE process (NVML queries made from the benchmark's own
    process now and then hold up its kernel launches for a millisecond or two, which a 7 ms timed region cannot hide).
    The samples carry wall-clock times; those inside [t_begin, t_end] are the ones reported."""
    REASONS = {
        0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
        0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
        0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting",
    }

    def __init__(self, index: int):
        import subprocess
        self.proc = None
        self.t_begin = self.t_end = None
        try:
            import torch
            props = torch.cuda.get_device_properties(index)
            bus = ""
            if hasattr(props, "pci_bus_id"):
                bus = "%02x:%02x.0" % (props.pci_bus_id, getattr(props, "pci_device_id", 0))
            self.proc = subprocess.Popen([sys.executable, "-c", _SAMPLER_SRC, bus, str(index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
        except Exception as e:  # pragma: no cover
            log(f"[bench] clock sampler unavailable ({e})")

    @property
    def active(self):
        return self.t_begin is not None and self.t_end is None

    @active.setter
    def active(self, on: bool):
        if on:
            self.t_begin, self.t_end = time.time(), None
        else:
            self.t_end = time.time()

    def start(self):
        pass

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:  # pragma: no cover
            self.proc.kill()
            out = ""
        sm_max, clocks, reasons = None, [], set()
        for line in out.splitlines():
            parts = line.split()
            if len(parts) == 2 and parts[0] == "max":
                sm_max = int(parts[1])
            elif len(parts) == 3:
                t, clk, mask = float(parts[0]), int(parts[1]), int(parts[2])
                if self.t_begin is not None and self.t_begin <= t <= (self.t_end or t):
                    clocks.append(clk)
                    for bit, name in self.REASONS.items():
                        if mask & bit and name != "gpu_idle":
                            reasons.add(name)
        s_ = sorted(clocks)
        return {"sm_mhz": s_[len(s_) // 2] if s_ else None, "sm_max_mhz": sm_max, "reasons": sorted(reasons),
                "samples": len(s_)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference algorithm on the host cores
# ------------------------------------------------------------------------------------------------
class CpuPort:
    """encode+decode of `n_items` x T frames of the bench workload with the CPU port (oracle/)."""

    def __init__(self, n_items: int):
        import oracle
        from tests.golden import gen
        self.oracle = oracle
        self.embed = gen.codebooks(900, CFG["K"], CFG["C"], CFG["D"])
        self.x = gen.normalish(901, (n_items, CFG["D"], CFG["T"]))
        oracle.set_threads(0)
        self.threads = min(oracle.get_threads(), 64)
        self.frames = n_items * CFG["T"]

    def one_pass(self) -> float:
        t0 = time.perf_counter()
        codes = self.oracle.encode(self.x, self.embed)
        q = self.oracle.decode(codes, self.embed)
        dt = time.perf_counter() - t0
        assert q.shape == self.x.shape
        return dt


class CpuReference:
    """encode+decode of the FULL bench batch (B x T frames) with the unmodified reference
    (audiocraft.quantization.ResidualVectorQuantizer from oracle/_ref, staged by oracle/fetch_ref.py) on the
    host cores: torch CPU, fp32, every host thread."""

    def __init__(self):
        import torch
        from oracle import ref_shim
        if not ref_shim.available():
            raise RuntimeError("no staged reference (oracle/_ref); run oracle/fetch_ref.py where /root/reference exists")
        q = ref_shim.load()
        self.torch = torch
        self.threads = os.cpu_count() or 1
        torch.set_num_threads(self.threads)
        torch.manual_seed(0)
        self.rvq = q.ResidualVectorQuantizer(dimension=CFG["D"], n_q=CFG["K"], bins=CFG["C"], kmeans_init=False).eval()
        self.x = torch.randn(CFG["B"], CFG["D"], CFG["T"], generator=torch.Generator().manual_seed(1234))
        self.frames = CFG["B"] * CFG["T"]
        self.root = ref_shim.REFERENCE_ROOT

    def one_pass(self) -> float:
        with self.torch.no_grad():
            t0 = time.perf_counter()
            codes = self.rvq.encode(self.x)
            q = self.rvq.decode(codes)
            dt = time.perf_counter() - t0
        assert q.shape == self.x.shape and codes.shape == (CFG["B"], CFG["K"], CFG["T"])
        return dt


def cpu_model() -> str:
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def make_cpu_arm():
    """(object with one_pass()/frames/threads, kind, sample text): the staged reference when it is there, else the port."""
    try:
        ref = CpuReference()
        return ref, "reference", (f"the full batch ({CFG['B']} x {CFG['T']} = {ref.frames} frames) per step, unmodified "
                                  f"audiocraft.quantization on torch CPU fp32, {ref.threads} threads, {cpu_model()}")
    except Exception as e:
        log(f"[bench] reference not staged ({e}); timing the C port instead")
        port = CpuPort(8)
        return port, "port", f"8 of {CFG['B']} batch items ({port.frames} frames) of the workload per step, C port, {cpu_model()}"


def run_reference_arm(args, rank: int):
    if rank != 0:
        return 0
    arm, kind, sample = make_cpu_arm()
    for _ in range(max(1, args.warmup)):
        arm.one_pass()
    times = [arm.one_pass() for _ in range(args.steps)]
    total = sum(times)
    value = arm.frames * len(times) / total / 1e6
    cpu = {"value": round(value, 6), "unit": UNIT, "cores": arm.threads, "kind": kind, "sample": sample}
    if kind == "reference":  # the C port beside it (second key; the port is the faster of the two)
        try:
            port = CpuPort(8)
            port.one_pass()
            dt = min(port.one_pass() for _ in range(3))
            cpu["port_value"] = round(port.frames / dt / 1e6, 6)
            cpu["port_sample"] = f"C port of the same algorithm (oracle/), 8 of {CFG['B']} items, {port.threads} threads, best of 3"
        except Exception as e:  # pragma: no cover
            log(f"[bench] port timing failed: {e}")
    line = {
        "impl": "reference", "metric": METRIC, "value": round(value, 6), "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(1e3 * total / len(times), 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": sample},
        "cpu_baseline": cpu,
        "e2e": {"value": round(value, 6), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def run_c5_share(dev, rank: int, world: int):
    """BASELINE.json configs[4]: 2^24 frames (B=8192 x T=2048), n_q=8, bins=2048, sharded by batch items over the
    ranks, codes all-gathered (int16 on the wire).  Every rank encodes + decodes its share; at N = 1 the share of an
    8-GPU job (2^21 frames) is timed, without a gather.  Informational extra key, not the bench value."""
    import torch
    import torch.distributed as dist

    from audiocraft_b200.quantization import ResidualVectorQuantizer
    from audiocraft_b200.sharded import ShardedRVQ
    K5, C5, D5, T5, B5_total = 8, 2048, 128, 2048, 8192
    ranks = world if world > 1 else 8
    b_loc = B5_total // ranks
    torch.manual_seed(5)
    m5 = ResidualVectorQuantizer(dimension=D5, n_q=K5, bins=C5, kmeans_init=False).to(dev).eval()
    m5.vq.check_codes = False
    sh = ShardedRVQ(m5) if world > 1 else None
    g = torch.Generator(device=dev).manual_seed(4321 + rank)
    x = torch.randn(b_loc, D5, T5, device=dev, generator=g)
    gshape = (B5_total, K5, T5)

    def one():
        codes = sh.encode_local(x, gshape) if sh else m5.encode(x)
        q = m5.decode(codes)
        ev = None
        if sh:
            _, ev = sh.gather_codes_async(codes, gshape)
        return q, ev

    with torch.no_grad():
        for _ in range(2):
            _, ev = one()
        if ev is not None:
            torch.cuda.current_stream(dev).wait_event(ev)
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 3
        a.record()
        for _ in range(reps):
            _, ev = one()
        if ev is not None:
            torch.cuda.current_stream(dev).wait_event(ev)
        b.record()
        torch.cuda.synchronize(dev)
        m5.vq.check_errors()
    ms = a.elapsed_time(b) / reps
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    frames = b_loc * T5
    flops = 2.0 * D5 * C5 * K5 * frames
    del x
    torch.cuda.empty_cache()
    return {"workload": "configs[4]: 2^24 frames n_q=8 bins=2048 dim=128 over %d ranks, encode+decode+all-gather of codes" % ranks
            if world > 1 else "configs[4]: one rank's share at 8 GPUs (2^21 frames), encode+decode, no gather",
            "frames_per_rank": frames, "ms_per_step": round(ms, 3),
            "mframes_per_s_aggregate": round(frames * (world if world > 1 else 1) / ms / 1e3, 2),
            "encode_decode_tflops_per_gpu": round(flops / ms / 1e9, 1)}



def run_gpu_arm(args, rank: int, world: int, local_rank: int):
    import torch
    import torch.distributed as dist

    from audiocraft_b200 import kernels
    from audiocraft_b200.quantization import ResidualVectorQuantizer

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    D, C, K, B, T = (CFG[k] for k in "DCKBT")
    frames = B * T
    torch.manual_seed(0)
    m = ResidualVectorQuantizer(dimension=D, n_q=K, bins=C, kmeans_init=False).to(dev).eval()
    m.vq.check_codes = False  # the in-kernel range check still runs; the host reads the word once, after timing
    if args.algo:
        m.vq.encode_algo = args.algo
    from audiocraft_b200.sharded import ShardedRVQ
    sharded = ShardedRVQ(m) if world > 1 else None  # broadcasts the codebooks from rank 0 once
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    xs = [torch.randn(B, D, T, device=dev, generator=g) for _ in range(N_ROTATING)]

    gshape = (B * world, K, T)  # every rank contributes one batch of configs[1]: the gathered codes are [B*world, K, T]

    def encode_step(x):
        # N > 1: the encode kernel also writes the codes as int16 into the send buffer of the all-gather
        return sharded.encode_local(x, gshape) if world > 1 else m.encode(x)

    class _NoEvent:
        def record(self):
            pass

    no_events = [_NoEvent()] * 3

    # The decode of a step runs on a second stream of lower priority: the next step's encode does not depend on it, so
    # its thread blocks fill the SMs that the encode kernel's CTAs leave as they finish (config 2's 750 tiles leave
    # 138 of 148 CTAs done a tenth of the kernel's time before the ten that own six tiles).  BENCH_SERIAL=1: one stream.
    overlap = not os.environ.get("BENCH_SERIAL")
    main_stream = torch.cuda.Stream(dev, priority=-1) if overlap else torch.cuda.current_stream(dev)
    dec_stream = torch.cuda.Stream(dev, priority=int(os.environ.get("BENCH_DEC_PRIO", "0"))) if overlap else None
    enc_done = [torch.cuda.Event() for _ in range(8)]
    # experiment (BENCH_ENC_STREAMS=2): consecutive encodes on two streams, so that the next kernel's CTAs start on the SMs
    # the previous kernel's CTAs have left
    n_enc = int(os.environ.get("BENCH_ENC_STREAMS", "1")) if overlap and world == 1 else 1  # (kernel times then overlap: not the default)
    enc_streams = [torch.cuda.Stream(dev, priority=-1) for _ in range(n_enc)] if n_enc > 1 else None

    def step(i, e3=no_events):
        """One step, the same code (and so the same allocation pattern) untimed and timed: encode, decode, and for
        N > 1 the all-gather of the codes on a side stream (the next batch does not depend on the gathered codes).
        Returns the event of the all-gather."""
        x = xs[i % N_ROTATING]
        if enc_streams is not None:
            es = enc_streams[i % n_enc]
            es.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(es):
                e3[0].record()
                codes = encode_step(x)
                e3[1].record()
                ed = enc_done[i % len(enc_done)]
                ed.record()
            codes.record_stream(es)
            dec_stream.wait_event(ed)
            with torch.cuda.stream(dec_stream):
                q = m.decode(codes)
                e3[2].record()
            codes.record_stream(dec_stream)
            return None
        e3[0].record()
        codes = encode_step(x)
        e3[1].record()
        if overlap:
            ed = enc_done[i % len(enc_done)]
            ed.record()
            dec_stream.wait_event(ed)
            with torch.cuda.stream(dec_stream):
                q = m.decode(codes)
                e3[2].record()
            codes.record_stream(dec_stream)
        else:
            q = m.decode(codes)
            e3[2].record()
        if world > 1:
            _, gathered = sharded.gather_codes_async(codes, gshape)  # int16 on the wire, widened by one kernel
            return gathered
        return None

    def sync_all(settle: bool = False):
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        if settle and not os.environ.get("BENCH_NO_GATE"):
            # (BENCH_NO_GATE=1: for runs under ncu, which serialises launches)
            # After an idle spell the GPUs of the ranks begin to execute a millisecond or two apart, and the code
            # exchange couples the ranks: whoever starts first then waits for the last one inside the timed region.
            # So every rank parks its stream behind a gate kernel (the GPU is awake, nothing else runs), the ranks
            # agree on a wall-clock time (one host, one clock), and each host opens its gate at that time.  The timed
            # region -- start event, K steps, end event -- is enqueued behind the gate.
            from audiocraft_b200 import _lib as rvq_lib
            gate = torch.zeros(1, dtype=torch.int32).pin_memory()
            rc = rvq_lib.load().rvq_debug_gate(gate.data_ptr(), int(1e9), torch.cuda.current_stream(dev).cuda_stream)
            rvq_lib.check(rc, "rvq_debug_gate")
            return gate
        return None

    def agree_on_opening_time():
        """Called when everything but the enqueueing of the steps is done: a wall-clock time 10 ms from the latest rank."""
        t0 = torch.tensor([time.time() + 0.01], dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t0, op=dist.ReduceOp.MAX, group=gate_group)
        return float(t0.item())

    gate_group = dist.new_group(backend="gloo") if world > 1 else None  # host-side agreement, no GPU work

    # ---- device-resident timing -------------------------------------------------------------
    with torch.no_grad(), torch.cuda.stream(main_stream):
        # the sampling thread is started before the warm-up so that its start-up cost is not in the timed region
        sampler = ClockSampler(local_rank) if rank == 0 and not os.environ.get("BENCH_NO_SAMPLER") else None
        if sampler:
            sampler.start()
        for i in range(args.warmup):
            step(i)
        sync_all()
        # W steps are a few milliseconds: keep stepping (untimed) until the GPU has been busy for about 60 ms, so
        # that a fresh box (idle clocks, cold L2 / TLBs, first-use work in the driver) is out of the numbers.
        # The same number of extra steps on every rank: the count is fixed from the measured step time of rank 0.
        t_w = time.perf_counter()
        for i in range(20):
            step(i)
        sync_all()
        per_step = max((time.perf_counter() - t_w) / 20, 1e-5)
        extra = torch.tensor([int(min(2000, max(0, args.pre_ms * 1e-3 / per_step)))], device=dev)
        if world > 1:
            dist.broadcast(extra, src=0)
        for i in range(int(extra.item())):
            step(i)
        sync_all()
        ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        # torch creates the CUDA event at its first record(): do that, and the interpreter's garbage collection, outside
        # of the timed region, so that the host keeps ahead of the GPU from the first timed step on
        for e3 in ev:
            for e in e3:
                e.record()
        start.record(); end.record()
        gate = sync_all(settle=True)
        import gc
        gc.collect()
        gc.disable()
        kernels.launch_count(reset=True)
        if sampler:
            sampler.active = True
        if sharded is not None and os.environ.get("BENCH_TRACE"):
            sharded.trace = []
        t_open = agree_on_opening_time() if gate is not None else None
        start.record()          # behind the gate when there is one
        host_t = [time.perf_counter()]
        for i in range(args.steps):
            gathered = step(i, ev[i])
            host_t.append(time.perf_counter())
            if gate is not None and i == min(1, args.steps - 1):
                # two steps are queued behind the gate: open it at the agreed time (the host keeps ahead of the GPU)
                while time.time() < t_open:
                    pass
                gate[0] = 1
                gate = None
        if world > 1:
            torch.cuda.current_stream(dev).wait_event(gathered)  # every all-gather ends inside the timed region
        if enc_streams is not None:
            for es in enc_streams:
                torch.cuda.current_stream(dev).wait_stream(es)
        if overlap:
            torch.cuda.current_stream(dev).wait_stream(dec_stream)  # ... and every decode
        end.record()
        sync_all()
        gc.enable()
        if sampler:
            sampler.active = False
        launches = kernels.launch_count()
        clocks = sampler.stop() if sampler else None
        m.vq.check_errors()
    elapsed_ms = start.elapsed_time(end)
    if sharded is not None and getattr(sharded, "trace", None):
        log("[bench] gather trace (start after timed start, push ms, widen ms): " + " ".join(
            f"{start.elapsed_time(a):.2f}/{a.elapsed_time(b):.2f}/{b.elapsed_time(c):.2f}" for a, b, c in sharded.trace))
    log("[bench] per-step host enqueue ms: " + " ".join(f"{(b - a) * 1e3:.3f}" for a, b in zip(host_t, host_t[1:])))
    log("[bench] per-step encode ms: " + " ".join(f"{e[0].elapsed_time(e[1]):.3f}" for e in ev))
    enc_ms = sum(e[0].elapsed_time(e[1]) for e in ev) / args.steps
    if overlap:
        # the decode kernel alone (in the steps it shares the GPU with the next encode): a serial loop after the timing
        with torch.no_grad():
            cs2 = [m.encode(xs[i]) for i in range(2)]
            for i in range(4):
                q = m.decode(cs2[i % 2])
            torch.cuda.synchronize(dev)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda._sleep(6_000_000)  # ~3 ms of GPU spin: the 20 launches are queued before the first one runs
            a.record()
            for i in range(20):
                q = m.decode(cs2[i % 2])
            b.record()
            torch.cuda.synchronize(dev)
            dec_ms = a.elapsed_time(b) / 20
            del cs2, q
    else:
        dec_ms = sum(e[1].elapsed_time(e[2]) for e in ev) / args.steps
    t = torch.tensor([elapsed_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    elapsed_ms = float(t.item())

    # ---- N > 1: are the gathered codes right?  (outside the timed region) ---------------------------------------
    gather_check = None
    if world > 1:
        with torch.no_grad():
            codes0 = encode_step(xs[0])
            full = sharded.gather_codes(codes0, gshape)          # [B*world, K, T] on every rank
            torch.cuda.synchronize(dev)
            if rank == 0:
                bad = 0
                for r in range(world):
                    gr = torch.Generator(device=dev).manual_seed(1234 + r)   # rank r's first input batch, regenerated here
                    xr = torch.randn(B, D, T, device=dev, generator=gr)
                    want = m.encode(xr)
                    bad += int((full[r * B:(r + 1) * B] != want).sum())
                gather_check = "ok" if bad == 0 else f"FAILED: {bad} codes differ"
                log(f"[bench] gather_check over {world} ranks: {gather_check}")
            sync_all()

    # ---- one rank's share of configs[4] (16M frames, n_q=8, strong scaling over the ranks), with its gather ----
    c5 = None
    if not args.skip_c5:
        c5 = run_c5_share(dev, rank, world)

    # ---- end to end through the public API with host buffers ---------------------------------
    # audiocraft_b200.host_pipeline.HostPipeline: pinned host latents in, pinned host codes + quantized latents
    # out, chunked over copy / compute / copy streams.  Every byte of every step crosses PCIe inside the timed region.
    from audiocraft_b200.host_pipeline import HostPipeline
    pipe = HostPipeline(m, dev, chunk_items=32)
    x_host = [torch.randn(B, D, T).pin_memory() for _ in range(2)]
    codes_host = [torch.empty((B, K, T), dtype=torch.int64).pin_memory() for _ in range(2)]
    q_host = [torch.empty((B, T, D), dtype=torch.float32).pin_memory() for _ in range(2)]

    with torch.no_grad():
        for i in range(max(20, args.warmup)):  # the first host<->device copies of a process are several times slower
            pipe.encode_decode(x_host[i % 2], codes_host[i % 2], q_host[i % 2])
        pipe.synchronize()
        sync_all()
        s2, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s2.record()
        last = None
        for i in range(args.steps):
            last = pipe.encode_decode(x_host[i % 2], codes_host[i % 2], q_host[i % 2])
        torch.cuda.current_stream(dev).wait_event(last)
        e2.record()
        pipe.synchronize()
        sync_all()
    e2e_ms = s2.elapsed_time(e2)
    t = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms = float(t.item())
    h2d = B * D * T * 4
    d2h = B * K * T * 8 + B * T * D * 4

    # ---- decode on configs[3] (B=256 x T=3000): the HBM-bound yardstick, informational ---------
    dec_c4 = None
    if rank == 0 and not args.skip_c4:
        with torch.no_grad():
            B4, T4 = 256, 3000
            cs = [torch.randint(0, C, (B4, K, T4), device=dev) for _ in range(2)]
            out = None
            for i in range(3):
                out = m.decode(cs[i % 2])
            torch.cuda.synchronize(dev)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 20
            a.record()
            for i in range(reps):
                out = m.decode(cs[i % 2])
            b.record()
            torch.cuda.synchronize(dev)
            m.vq.check_errors()
            us = a.elapsed_time(b) * 1e3 / reps
            nbytes = B4 * T4 * (8 * K + 4 * D)
            dec_c4 = {"workload": "configs[3]: decode-only B=256 x T=3000 n_q=4", "us_per_launch": round(us, 2),
                      "gframes_per_s": round(B4 * T4 / us / 1e3, 3), "achieved_gbs": round(nbytes / us / 1e3, 1)}
            del cs, out

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    exchange_kind = "none"
    if sharded is not None:
        exchange_kind = ("peer stores into symmetric memory + flags, librvq_b200.so kernels"
                         if any(v is not None for v in sharded._p2p.values()) else "NCCL all_gather + one widening kernel")
    peaks = measured_peaks()
    total_frames = frames * world * args.steps
    value = total_frames / (elapsed_ms * 1e-3) / 1e6
    enc_flops = 2.0 * D * C * K * frames  # algorithmic: the ||x||^2 - 2 x.c + ||c||^2 contraction only
    achieved_tf = enc_flops / (enc_ms * 1e-3) / 1e12
    traffic = None
    try:  # DRAM bytes per launch of the encode kernel from the committed ncu --set full capture (profiles/)
        with open(os.path.join(ROOT, "profiles", "r3_traffic.json")) as f:
            traffic = json.load(f).get("rvq_encode_tc_kernel_config2_bytes")
    except Exception:
        pass
    roofline = {"bound": "tensor", "kernel": "rvq_encode (all stages fused)", "achieved": round(achieved_tf, 3),
                "peak": peaks["bf16"], "unit": "TFLOP/s", "frac": round(achieved_tf / peaks["bf16"], 5),
                "traffic": traffic, "peak_source": peaks["source"] + ", dense bf16 burst",
                "us_per_launch": round(enc_ms * 1e3, 2)}
    if peaks.get("bf16_sustained"):
        # informational: under back-to-back steps the GPU may sit at its power cap (see "clocks"), the regime of the
        # driver's sustained figure; `frac` stays on the stricter burst peak
        roofline["peak_sustained"] = peaks["bf16_sustained"]
        roofline["frac_of_sustained"] = round(achieved_tf / peaks["bf16_sustained"], 5)
    if dec_c4 is not None:
        dec_c4["peak_gbs"] = peaks["hbm"]
        dec_c4["frac"] = round(dec_c4["achieved_gbs"] / peaks["hbm"], 4)

    cpu = None
    if world == 1 and not args.skip_cpu:
        arm, kind, sample = make_cpu_arm()
        arm.one_pass()
        dt = min(arm.one_pass() for _ in range(3 if kind == "reference" else 5))
        cpu = {"value": round(arm.frames / dt / 1e6, 6), "unit": UNIT, "cores": arm.threads, "kind": kind,
               "sample": sample + ", encode+decode, best of " + ("3" if kind == "reference" else "5")}
        if kind == "reference":
            port = CpuPort(8)
            port.one_pass()
            dtp = min(port.one_pass() for _ in range(3))
            cpu["port_value"] = round(port.frames / dtp / 1e6, 6)
            cpu["port_cores"] = port.threads

    line = {
        "metric": METRIC, "value": round(value, 4), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(elapsed_ms / args.steps, 4), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f16 tensor-core search + f32 exact re-score (codes bit-exact)",
        "data": "synthetic",
        "config": {"workload": WORKLOAD, "frames_per_gpu_per_step": frames,
                   "l2": f"{N_ROTATING} rotating input batches ({N_ROTATING * h2d // 1000000} MB) > 126 MB L2",
                   "encode_algo": m.vq.encode_algo,
                   "untimed_steps_before_timing": args.warmup + 20 + int(extra.item()),
                   "start": "timed region enqueued behind a gate kernel opened at an agreed wall-clock time",
                   "streams": ("encode on a high-priority stream, the decode of every step on a second stream (it fills the "
                               "SMs the encode kernel's CTAs leave as they finish); all complete inside the timed region")
                   if overlap else "one stream",
                   "collective": ("codes of every step exchanged as int16 (" + exchange_kind + "), side stream, all complete "
                                  "inside the timed region") if world > 1 else "none"},
        "roofline": roofline,
        "cpu_baseline": cpu,
        "e2e": {"value": round(frames * world * args.steps / (e2e_ms * 1e-3) / 1e6, 4), "unit": UNIT,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "kernels": {"encode_ms": round(enc_ms, 4), "decode_ms": round(dec_ms, 4),
                    "encode_mframes_per_s": round(frames / enc_ms / 1e3, 3),
                    "decode_mframes_per_s": round(frames / dec_ms / 1e3, 3)},
        "decode_c4": dec_c4,
        "gather_check": gather_check,
        "c5_share": c5,
    }
    emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--algo", default="", choices=["", "auto", "exact", "tc"])
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--pre-ms", type=float, default=60.0, help="untimed steps before the timed region, in ms of GPU time")
    ap.add_argument("--skip-c4", action="store_true")
    ap.add_argument("--skip-c5", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    claim_stdout()
    if args.impl == "reference":
        return run_reference_arm(args, rank)
    if world != args.gpus:
        log(f"[bench] --gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run for N > 1")
    return run_gpu_arm(args, rank, world, local_rank)


if __name__ == "__main__":
    sys.exit(main())
