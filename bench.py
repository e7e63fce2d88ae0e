#!/usr/bin/env python
"""bench.py -- Compress/Decompress MCells/s of the CvxCompress hot path on B200 (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W [--config C]   # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K ...      # the reference's own OpenMP/AVX CPU path (oracle/_ref)

A "step" = one Compress of the volume followed by one Decompress of the stream it produced
(CvxCompress.cpp:231-427 and :463-571).  value = cells / (t_compress + t_decompress) in MCells/s;
the two directions are also reported on their own.

Workload (default --config 3): BASELINE config 3 -- nx,ny,nz = 1056,1248,960 fp32 (5.06 GB), 32^3 blocks,
scale 1e-2, global RMS, G2 seismic-like synthetic data.  With N>1 ranks this ONE volume is sharded
by z-slabs of whole blocks -- strong scaling, what the reference does with its workers
(CvxCompress.cpp:318, :528) -- through cvxgpu_compress_sharded / cvxgpu_decompress_sharded (C ABI + NCCL), each
direction replayed from a CUDA graph.  `--scaling weak` gives every rank a 960-plane slab of a 960*N-plane volume; a
short weak run is also reported inside the strong line ("weak").  The other BASELINE configs: --config 1 (352x416x320 G1),
2 (wavelet fwd+inv sweep 8^3..256^3), 4 (1024^3, 16x16x64 blocks, local RMS), 5 (1280^3 decompress-only of a stream
written by the reference on the host cores).

Timing: CUDA events on the launching stream, barrier + synchronize on both sides, max over ranks.
L2: the volume (5 GB) and the decoded output (5 GB) are each 40x larger than the 126 MB L2.
"""
import argparse
import ctypes as C
import json
import math
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

UNIT = "MCells/s"
SCALE = 1e-2
# BASELINE.json configs; nx is the fastest axis ("960x1248x1056" in BASELINE.json is slow x mid x fast)
CONFIGS = {
    1: dict(nx=352, ny=416, nz=320, b=(32, 32, 32), local=False, data="g1",
            name="BASELINE config 1: Test_With_Generated_Input 352x416x320 (nx,ny,nz) fp32, 32^3 blocks, global RMS"),
    3: dict(nx=1056, ny=1248, nz=960, b=(32, 32, 32), local=False, data="g2",
            name="BASELINE config 3: Generated 1056x1248x960 (nx,ny,nz) fp32 volume, 32^3 blocks, global RMS"),
    4: dict(nx=1024, ny=1024, nz=1024, b=(16, 16, 64), local=True, data="g2",
            name="BASELINE config 4: SEAM-Basin-scale synthetic 1024^3 fp32 volume (4.29 GB), 16x16x64 blocks, local RMS"),
    5: dict(nx=1280, ny=1280, nz=1280, b=(32, 32, 32), local=False, data="g2",
            name="BASELINE config 5: decompress-only round trip of a 1280^3 fp32 volume (8.39 GB), 32^3 blocks, stream written by the reference"),
}
NX = NY = NZ = BX = BY = BZ = 0        # set by select_config()
LOCAL = False
SAMPLE_NZ = 320                       # CPU legs: the first 320 planes of the same volume


def select_config(args):
    global NX, NY, NZ, BX, BY, BZ, LOCAL, SAMPLE_NZ
    c = CONFIGS[args.config if args.config in CONFIGS else 3]
    NX, NY, NZ = c["nx"], c["ny"], c["nz"]
    BX, BY, BZ = c["b"]
    LOCAL = c["local"]
    if args.nz:
        NZ = args.nz
    if args.data is None:
        args.data = c["data"]
    SAMPLE_NZ = min(320, NZ)
    if BZ > 32:
        SAMPLE_NZ = min(NZ, 5 * BZ)
    return c


def metric_name(args, world):
    if args.config == 5:
        return "Decompress MCells/s of a reference-written stream, %dx%dx%d fp32, %d^3 blocks, %d B200" % (NX, NY, NZ, BX, world)
    return "Compress+Decompress round-trip MCells/s, %dx%dx%d fp32, %dx%dx%d blocks, %s RMS, %s arithmetic" % (
        NX, NY, NZ if args.scaling == "strong" else NZ * world, BX, BY, BZ, "local" if LOCAL else "global", args.transform.upper())


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", type=int, default=3, choices=[1, 2, 3, 4, 5], help="BASELINE.json config (default 3, the metric's)")
    ap.add_argument("--transform", default=os.environ.get("CVX_BENCH_TRANSFORM", "lifting"), choices=["exact", "lifting"],
                    help="lifting (default): FMA lifting arithmetic, coefficients within 1e-6 of block magnitude, SNR within 0.1 dB; "
                         "exact: the reference's operation order, bit-identical stream (also timed and reported as exact_mode)")
    ap.add_argument("--scaling", default="strong", choices=["weak", "strong"],
                    help="N>1: strong (default) = the one volume of the config sharded over the ranks; weak = one such volume per rank")
    ap.add_argument("--data", default=None, choices=["g1", "g2"])
    ap.add_argument("--nz", type=int, default=0, help="planes per volume (smaller for quick checks)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer end-to-end leg")
    ap.add_argument("--no-graph", action="store_true", help="N>1: enqueue every step eagerly instead of replaying CUDA graphs")
    return ap.parse_args()


# ---- clocks during the timed region (B200_PROFILING.md recipe; NVML polled from a thread) -----------------
class ClockSampler:
    """Polls SM clock, power and the clock-event (throttle) reasons of one GPU every ~2 ms through NVML while the
    timed region runs (the region lasts tens of ms, too short for `nvidia-smi -lms`)."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown"}

    def __init__(self, torch_index):
        self.ok = False
        self.sm, self.pw, self.bits = [], [], 0
        self.stop_flag = False
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            self.nv = pynvml
            try:
                uuid = str(torch.cuda.get_device_properties(torch_index).uuid)
                if not uuid.startswith("GPU-"):
                    uuid = "GPU-" + uuid
                self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(torch_index)
            self.max_sm = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception as e:          # no NVML: report that, do not fail the bench
            self.err = repr(e)

    def _poll(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                self.pw.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
                self.bits |= int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        if self.ok:
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()

    def stop(self):
        if not self.ok:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["NVML unavailable: " + getattr(self, "err", "?")]}
        self.stop_flag = True
        self.thread.join(timeout=2)
        sm = sorted(self.sm)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_sm,
                "power_w_max": max(self.pw) if self.pw else None, "samples": len(sm),
                "reasons": sorted(n for b, n in self.REASONS.items() if self.bits & b)}


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---- the reference's own CPU implementation (oracle/_ref, else the oracle port) ---------------------------
def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def make_host_sample(data, nz_total, sample_nz):
    """First `sample_nz` planes of the workload volume, on the host (numpy [z, y, x])."""
    import numpy as np
    from cvxcompress_b200 import synth
    try:
        import torch
        if torch.cuda.is_available():
            gen = synth.g1_torch if data == "g1" else synth.g2_torch
            return gen(NX, NY, sample_nz, "cuda:0", z0=0, nz_total=nz_total).cpu().numpy()
    except Exception:
        pass
    if data == "g1":
        return synth.g1(NX, NY, sample_nz)
    out = np.empty((sample_nz, NY, NX), np.float32)
    for k in range(sample_nz):                       # plane by plane: bounded memory
        out[k] = synth.g2_slab(NX, NY, 1, z0=k, nz_total=nz_total)[0]
    return out


class CpuPath:
    """Compress/Decompress of the reference on the host cores.  kind 'reference' = the unmodified
    reference build (oracle/_ref/libcvxcompress_ref.so, OpenMP over blocks, all cores);
    kind 'port' = the single-threaded C restatement (oracle/libcvx_oracle.so)."""

    def __init__(self):
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_libs as ol
        self.ol = ol
        self.kind = "reference" if ol.have_ref() else "port"
        self.cores = host_cores() if self.kind == "reference" else 1

    def prepare(self, vol):
        import numpy as np
        self.vol = np.ascontiguousarray(vol).reshape(-1)
        self.nz, self.ny, self.nx = vol.shape
        cap = self.ol.oracle().cvxo_max_compressed_bytes(self.nx, self.ny, self.nz, BX, BY, BZ, int(LOCAL)) // 4 + 4096
        self.stream = self.ol.aligned_u8(cap)
        self.out = np.zeros(self.vol.size, dtype=np.float32)
        self.length = C.c_long(0)

    def compress(self):
        t0 = time.perf_counter()
        if self.kind == "reference":
            self.ol.ref().cvx_compress_th(SCALE, self.vol, self.nx, self.ny, self.nz, BX, BY, BZ, bool(LOCAL), self.stream,
                                          self.cores, C.byref(self.length))
        else:
            self.ol.oracle().cvxo_compress(SCALE, self.vol, self.nx, self.ny, self.nz, BX, BY, BZ, int(LOCAL), self.stream,
                                           C.byref(self.length))
        return time.perf_counter() - t0

    def decompress(self):
        t0 = time.perf_counter()
        if self.kind == "reference":
            self.ol.ref().cvx_decompress_inplace_th(self.out, self.nx, self.ny, self.nz, self.stream, self.cores,
                                                    self.length.value)
        else:
            self.ol.oracle().cvxo_decompress(self.out, self.nx, self.ny, self.nz, self.stream)
        return time.perf_counter() - t0


def quality_np(vol, out):
    import numpy as np
    num = den = 0.0
    v = vol.reshape(-1); w = out.reshape(-1)
    step = 1 << 24
    for i in range(0, v.size, step):                  # chunked: no 13 GB fp64 temporaries
        a = v[i:i + step].astype(np.float64); b = w[i:i + step].astype(np.float64)
        num += float(np.sum((a - b) ** 2)); den += float(np.sum(a * a))
    err = math.sqrt(num / den) if den > 0 else 0.0
    return err, (-20.0 * math.log10(err) if err > 0 else float("inf"))


def cpu_leg(data, nz_total, steps, warmup, ours_on_sample=None):
    """Times `steps` round trips of the reference on the bounded sample; returns (dict for cpu_baseline, tc, td).
    ours_on_sample(vol_numpy) -> (compressed_length, snr_db) runs THIS repo's CUDA path on the same sample so that the
    two arms' ratio and SNR stand side by side (north_star: ratio within 0.1 %, SNR within 0.1 dB)."""
    sample_nz = min(SAMPLE_NZ, nz_total)
    cpu = CpuPath()
    if cpu.kind == "port":
        sample_nz = min(sample_nz, 64)               # scalar port: keep it to seconds
    vol = make_host_sample(data, nz_total, sample_nz)
    cpu.prepare(vol)
    for _ in range(warmup):
        cpu.compress(); cpu.decompress()
    tc = td = 0.0
    for _ in range(steps):
        tc += cpu.compress()
        td += cpu.decompress()
    tc /= steps; td /= steps
    cells = vol.size
    err, snr = quality_np(vol, cpu.out)
    info = {"value": cells / (tc + td) / 1e6, "unit": UNIT, "cores": cpu.cores, "kind": cpu.kind,
            "compress_mcells_s": cells / tc / 1e6, "decompress_mcells_s": cells / td / 1e6,
            "sample": "first %d z-planes of the %s volume (%dx%dx%d, %.1f MCells) compressed as a volume of its own, %d round trips "
                      "after %d warm-up, threads=%d" % (sample_nz, data.upper(), NX, NY, sample_nz, cells / 1e6, steps, warmup, cpu.cores),
            "compressed_length": int(cpu.length.value), "ratio": 4.0 * cells / cpu.length.value,
            "roundtrip_error": err, "snr_db": snr}
    if ours_on_sample is not None:
        ln, snr_o = ours_on_sample(vol)
        info["gpu_on_same_sample"] = {"compressed_length": int(ln), "ratio": 4.0 * cells / ln, "snr_db": snr_o,
                                      "ratio_rel_diff": (cpu.length.value - ln) / float(cpu.length.value),
                                      "snr_diff_db": snr_o - snr}
    return info, tc, td


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if args.config == 2:
        return run_reference_wavelet(args)
    info, tc, td = cpu_leg(args.data, NZ, args.steps, max(args.warmup, 1))
    if args.config == 5:                              # decompress-only metric
        value, ms = info["decompress_mcells_s"], td * 1e3
    else:
        value, ms = info["value"], (tc + td) * 1e3
    cfg = workload_config(args, 1)
    cfg["workload"] += "; THIS ARM: the reference's CPU path on a bounded sample -- " + info["sample"]
    line = {"impl": "reference", "metric": metric_name(args, 1), "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": cfg, "cpu_baseline": info,
            "compress_mcells_s": info["compress_mcells_s"], "decompress_mcells_s": info["decompress_mcells_s"],
            "ratio": info["ratio"], "snr_db": info["snr_db"],
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def workload_config(args, world):
    nz_total = NZ * world if args.scaling == "weak" else NZ
    nbz = -(-nz_total // BZ)
    most = -(-nbz // world)
    return {"workload": "%s, scale 1e-2, %s synthetic data" % (CONFIGS[args.config]["name"] if args.config in CONFIGS else "", args.data.upper())
                        + ("" if nz_total == CONFIGS.get(args.config, {}).get("nz", nz_total) else " [nz = %d]" % nz_total),
            "config": args.config, "nx": NX, "ny": NY, "nz": nz_total, "bx": BX, "by": BY, "bz": BZ, "scale": SCALE, "use_local_rms": bool(LOCAL),
            "transform": args.transform,
            "sharding": "z-slabs of whole blocks, %d rank(s): %d block rows, at most %d per rank" % (world, nbz, most),
            "slab_imbalance_ceiling": nbz / float(world * most),
            "l2": "inputs larger than L2 (volume and output per rank >> 126 MB L2); no flush needed"}


def numa_bind(local_rank):
    """Pin this process (and the pinned host buffers it first-touches) to the NUMA node of its GPU: with all ranks' copy
    threads and staging on node 0 the per-rank PCIe rate collapsed from 54 to 14 GB/s at 8 ranks (SCALE_r01)."""
    try:
        import torch
        bus = torch.cuda.get_device_properties(local_rank).pci_bus_id
        dom = getattr(torch.cuda.get_device_properties(local_rank), "pci_domain_id", 0)
        dev = torch.cuda.get_device_properties(local_rank).pci_device_id
        path = "/sys/bus/pci/devices/%04x:%02x:%02x.0/numa_node" % (dom, bus, dev)
        node = int(open(path).read().strip())
        if node < 0:
            return {"numa_node": node, "note": "the box exposes no NUMA topology for this GPU; nothing bound"}
        cpus = []
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus += list(range(int(lo), int(hi or lo) + 1))
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
        if allowed:
            os.sched_setaffinity(0, allowed)
        return {"numa_node": node, "cpus": len(allowed)}
    except Exception as e:
        return {"numa_node": None, "why": repr(e)}


# ---- this repo's CUDA path --------------------------------------------------------------------------------
def init_dist():
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the hot path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa = numa_bind(local_rank) if world > 1 else None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    return rank, world, local_rank, dev, numa


def time_kernel(fn, reps):
    import torch
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


class RoundTrip:
    """Compress + Decompress of this rank's share of one nx*ny*nz_total volume, device-resident, as two closures.
    world == 1: the whole-volume C calls (cvxgpu_compress_device_async / cvxgpu_decompress_device).
    world  > 1: cvxgpu_compress_sharded / cvxgpu_decompress_sharded (kernels + two ncclAllGather on the stream, no host
    synchronisation), each direction captured once in a CUDA graph and replayed."""

    def __init__(self, args, rank, world, dev, nz_total, use_graph=True):
        import torch
        from cvxcompress_b200 import api, sharded, synth
        self.torch, self.api, self.sharded = torch, api, sharded
        self.world, self.rank, self.dev, self.nz_total = world, rank, dev, nz_total
        self.g = sharded.Geometry(NX, NY, nz_total, BX, BY, BZ, rank, world)
        gen = synth.g1_torch if args.data == "g1" else synth.g2_torch
        self.slab = gen(NX, NY, self.g.nz_slab, dev, z0=self.g.z0, nz_total=nz_total)
        self.out_slab = torch.empty_like(self.slab)
        self.cells_rank = self.slab.numel()
        self.cells_total = NX * NY * nz_total
        self.graphs = None
        g = self.g
        tables = 8 * g.nblk + (4 * g.nblk if LOCAL else 0)
        if world == 1:
            # config 3 compresses 40:1; a third of the raw volume (+ tables) is ample, and the status word is checked
            cap = max(self.slab.numel() * 4 // 3, 1 << 20) + tables + 4096
            self.stream_buf = torch.empty(cap + 64, dtype=torch.uint8, device=dev)
            self.d_len = torch.zeros(1, dtype=torch.int64, device=dev)
            self.d_status = torch.zeros(1, dtype=torch.int32, device=dev)
            self.cap = cap
            self.compress()
            torch.cuda.synchronize()
            assert int(self.d_status.item()) == 0, "stream buffer too small"
            self.length = int(self.d_len.item())
        else:
            self.payload = torch.empty(max(self.slab.numel() * 4 // 3, 1 << 20) + 4096, dtype=torch.uint8, device=dev)
            self.ss = sharded.compress_slab(self.slab, NX, NY, nz_total, BX, BY, BZ, SCALE, LOCAL, payload=self.payload)
            torch.cuda.synchronize()
            self.ss.check()
            self.length = self.ss.total_length()
            self.decompress()
            torch.cuda.synchronize()
            if use_graph:
                # one eager round trip has sized every workspace; now record each direction once
                st = torch.cuda.Stream()
                st.wait_stream(torch.cuda.current_stream())
                gc, gd = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
                with torch.cuda.stream(st):
                    with torch.cuda.graph(gc, stream=st):
                        self._compress_eager()
                    with torch.cuda.graph(gd, stream=st):
                        self._decompress_eager()
                torch.cuda.current_stream().wait_stream(st)
                self.graphs = (gc, gd)

    def _compress_eager(self):
        if self.world == 1:
            self.api.compress_device_async(self.slab, BX, BY, BZ, SCALE, LOCAL, self.stream_buf[:self.cap], self.d_len, self.d_status)
        else:
            self.ss = self.sharded.compress_slab(self.slab, NX, NY, self.nz_total, BX, BY, BZ, SCALE, LOCAL, out=self.ss)

    def _decompress_eager(self):
        if self.world == 1:
            self.api.decompress_device(self.stream_buf, self.length, NX, NY, self.nz_total, out=self.out_slab)
        else:
            self.sharded.decompress_slab(self.ss, out=self.out_slab)

    def compress(self):
        if self.graphs:
            self.graphs[0].replay()
        else:
            self._compress_eager()

    def decompress(self):
        if self.graphs:
            self.graphs[1].replay()
        else:
            self._decompress_eager()

    def error(self):
        torch = self.torch
        num = torch.sum((self.slab.double() - self.out_slab.double()) ** 2)
        den = torch.sum(self.slab.double() ** 2)
        if self.world > 1:
            import torch.distributed as dist
            nd = torch.stack([num, den]); dist.all_reduce(nd); num, den = nd[0], nd[1]
        return float(torch.sqrt(num / den).item())

    def timed(self, K, W, barrier, sampler=None):
        """(ms_per_step, compress_ms, decompress_ms, host_wall_ms_per_step), max over ranks."""
        torch = self.torch
        for _ in range(W):
            self.compress(); self.decompress()
        ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(K)]
        barrier()
        if sampler:
            sampler.start()
        t0 = time.perf_counter()
        for k in range(K):
            ev[k][0].record(); self.compress(); ev[k][1].record(); self.decompress(); ev[k][2].record()
        torch.cuda.synchronize()
        t_host = time.perf_counter() - t0
        self.clocks = sampler.stop() if sampler else None
        barrier()
        tm = torch.tensor([ev[0][0].elapsed_time(ev[K - 1][2]) / K, sum(e[0].elapsed_time(e[1]) for e in ev) / K,
                           sum(e[1].elapsed_time(e[2]) for e in ev) / K, t_host * 1e3 / K], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            import torch.distributed as dist
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        return tuple(float(v) for v in tm.tolist())


def run_ours(args):
    if args.config == 2:
        return run_wavelet(args)
    if args.config == 5:
        return run_config5(args)
    import torch
    rank, world, local_rank, dev, numa = init_dist()
    import torch.distributed as dist
    from cvxcompress_b200 import _lib, api, sharded, synth
    L = _lib.lib()
    api.set_transform(args.transform)
    st = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)
    K, W = args.steps, args.warmup

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    nz_total = NZ * world if args.scaling == "weak" else NZ
    rt = RoundTrip(args, rank, world, dev, nz_total, use_graph=not args.no_graph)
    g, slab, out_slab = rt.g, rt.slab, rt.out_slab
    cells_rank, cells_total, length = rt.cells_rank, rt.cells_total, rt.length
    rt.decompress()
    torch.cuda.synchronize()

    # quality of the round trip (also proves the timed work is the real work): checked here and again after the last timed
    # decode.  The hard gate is the reference arm's SNR on the same data (cpu_leg, below); without it (N>1, --no-cpu) the
    # quantiser's own bound stays as a sanity check.
    def check_quality(e, where):
        if not (math.isfinite(e) and e < SCALE):
            sys.stderr.write("bench.py: INVALID RUN: round-trip error %r %s (bound %g)\n" % (e, where, SCALE))
            sys.exit(3)

    err = rt.error()
    check_quality(err, "after the first Decompress")
    snr = -20.0 * math.log10(err) if err > 0 else float("inf")

    sampler = ClockSampler(local_rank)
    ms_per_step, tc_ms, td_ms, host_ms = rt.timed(K, W, barrier, sampler)
    clocks = rt.clocks
    value = cells_total / (ms_per_step * 1e-3) / 1e6

    # -- per-kernel timings for the roofline: the hot kernels on their own, CUDA events on the launching stream --
    peak, peak_src = measured_peak_gbs()
    d_sumsq = torch.zeros(1, dtype=torch.float64, device=dev)
    d_rows = torch.zeros(max(g.b1 - g.b0, 1), dtype=torch.float64, device=dev)
    offs_t = torch.zeros(max(g.nblk_slab, 1), dtype=torch.int64, device=dev)
    bmf_t = torch.zeros(max(g.nblk_slab, 1), dtype=torch.float32, device=dev) if LOCAL else None
    pay_t = torch.empty(max(slab.numel() * 4 // 3, 1 << 20) + 4096, dtype=torch.uint8, device=dev)
    nbytes_t = torch.zeros(1, dtype=torch.int64, device=dev)
    stat_t = torch.zeros(1, dtype=torch.int32, device=dev)
    bmf_p = bmf_t.data_ptr() if LOCAL else None

    def k_sumsq():
        _lib.check(L.cvxgpu_sumsq_rows_device(None, slab.data_ptr(), NX * NY, BZ, g.nz_slab, d_rows.data_ptr(), st()))

    def k_encode():
        _lib.check(L.cvxgpu_encode_blocks_device(None, slab.data_ptr(), NX, NY, g.nz_slab, BX, BY, BZ, SCALE, int(LOCAL),
                                                 d_sumsq.data_ptr(), slab.numel(), offs_t.data_ptr(), bmf_p, pay_t.data_ptr(),
                                                 pay_t.numel(), nbytes_t.data_ptr(), stat_t.data_ptr(), st()))

    k_sumsq()
    _lib.check(L.cvxgpu_sum_rows_device(d_rows.data_ptr(), g.b1 - g.b0, d_sumsq.data_ptr(), st()))
    k_encode()
    torch.cuda.synchronize()
    assert int(stat_t.item()) == 0
    pbytes = int(nbytes_t.item())
    gm = 1.0 if LOCAL else 1.0 / (math.sqrt(float(d_sumsq.item()) / slab.numel()) * SCALE)

    def k_decode():
        _lib.check(L.cvxgpu_decode_blocks_device(None, out_slab.data_ptr(), NX, NY, g.nz_slab, BX, BY, BZ, gm, int(LOCAL),
                                                 offs_t.data_ptr(), bmf_p, pay_t.data_ptr(), pbytes, st()))

    reps = max(3, min(K, 20))
    stream_slab_bytes = pbytes + (12 if LOCAL else 8) * g.nblk_slab
    kern = {}
    is32 = (BX, BY, BZ) == (32, 32, 32)
    if is32:
        # LIFTING dispatches the warp-specialised kernels, EXACT the fused ones (cvx_kernels.cu launch_encode / launch_decode)
        enc_name, dec_name = ("k_encode32ws", "k_decode32ws") if args.transform == "lifting" else ("k_encode32", "k_decode32")
    elif max(BX, BY, BZ) <= 64 and 8192 <= BX * BY * BZ <= 32768:
        # config 4's 16 x 16 x 64 and its siblings: sparse list coder / windowed parser (cvx_encode_sparse.inl, cvx_decode_win.inl)
        enc_name, dec_name = "k_encode_sparse", "k_decode_win"
    else:
        enc_name, dec_name = "k_encode_fused", "k_decode_fused"
    klist = [] if LOCAL else [("k_sumsq_rows", k_sumsq, 4 * cells_rank)]
    klist += [(enc_name, k_encode, 4 * cells_rank + stream_slab_bytes), (dec_name, k_decode, 4 * cells_rank + stream_slab_bytes)]
    for name, fn, abytes in klist:
        ms = time_kernel(fn, reps)
        gbs = abytes / (ms * 1e-3) / 1e9
        kern[name] = {"ms": ms, "algorithmic_bytes": abytes, "achieved_gbs": gbs, "frac": gbs / peak,
                      "mcells_s": cells_rank / (ms * 1e-3) / 1e6}
    dom = max(kern, key=lambda k: kern[k]["ms"])
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")       # dram bytes per launch from the committed ncu --set full capture
    if os.path.exists(tp):
        try:
            with open(tp) as f:
                tj = json.load(f).get(args.transform, {}).get(dom)
                if tj:                                         # scale the capture's per-launch DRAM bytes to this launch's cell count
                    traffic = int(tj["dram_bytes"] * (cells_rank / float(tj["cells"])))
        except Exception:
            traffic = None
    cbytes = (4 if LOCAL else 8) * cells_rank + stream_slab_bytes
    dbytes = 4 * cells_rank + stream_slab_bytes
    roofline = {"bound": "hbm", "kernel": dom, "achieved": kern[dom]["achieved_gbs"], "peak": peak, "unit": "GB/s",
                "frac": kern[dom]["frac"], "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": kern[dom]["algorithmic_bytes"], "ms_per_launch": kern[dom]["ms"],
                "kernels": kern,
                "whole_call": {"compress_gbs": cbytes / (tc_ms * 1e-3) / 1e9, "decompress_gbs": dbytes / (td_ms * 1e-3) / 1e9,
                               "compress_frac": cbytes / (tc_ms * 1e-3) / 1e9 / peak,
                               "decompress_frac": dbytes / (td_ms * 1e-3) / 1e9 / peak,
                               "compress_frac_of_8TBs": cbytes / (tc_ms * 1e-3) / 1e9 / 8000.0,
                               "decompress_frac_of_8TBs": dbytes / (td_ms * 1e-3) / 1e9 / 8000.0}}
    err_end = rt.error()                            # out_slab now holds the last timed k_decode launch's output
    check_quality(err_end, "after the timed region")
    err = max(err, err_end)

    # -- end to end through the reference-facing call with HOST buffers, copies inside the timed region --
    e2e = e2e_pageable = None
    if not args.no_e2e:
        e2e, e2e_pageable = e2e_legs(args, rt, L, barrier, K)

    # -- the other arithmetic mode, device-resident, a few steps: EXACT = bit-identical stream, LIFTING = fast --
    exact_info = None
    if world == 1 and is32:
        other = "exact" if args.transform == "lifting" else "lifting"
        api.set_transform(other)
        rt.compress(); torch.cuda.synchronize()
        len_o = int(rt.d_len.item())
        def dec_o():
            api.decompress_device(rt.stream_buf, len_o, NX, NY, nz_total, out=out_slab)
        dec_o(); torch.cuda.synchronize()
        err_o = rt.error()
        tco = time_kernel(rt.compress, 3)
        tdo = time_kernel(dec_o, 3)
        eo = time_kernel(k_encode, 3)
        exact_info = {"transform": other, "compress_ms": tco, "decompress_ms": tdo, "compressed_length": len_o,
                      "snr_db": -20.0 * math.log10(err_o) if err_o > 0 else None,
                      "value": cells_total / ((tco + tdo) * 1e-3) / 1e6, "unit": UNIT,
                      "encode_kernel_ms": eo, "encode_kernel_frac": (4 * cells_rank + stream_slab_bytes) / (eo * 1e-3) / 1e9 / peak,
                      "decode_kernel_frac": (4 * cells_rank + stream_slab_bytes) / (tdo * 1e-3) / 1e9 / peak}
        api.set_transform(args.transform)

    # -- what limits strong scaling: the fixed cost of a sharded call (kernel launches inside the graph, the two 8-byte..
    #    ncclAllGather, the tail of a persistent grid), measured on a volume of one tiny block row per rank where the
    #    kernels themselves take a few microseconds --
    fixed = None
    if world > 1:
        tiny = torch.zeros((BZ, 64, 64), dtype=torch.float32, device=dev) + float(rank + 1)
        tss = sharded.compress_slab(tiny, 64, 64, BZ * world, BX, BY, BZ, SCALE, LOCAL)
        tout = torch.empty_like(tiny)
        sharded.decompress_slab(tss, out=tout)
        torch.cuda.synchronize()
        gt = torch.cuda.CUDAGraph()
        sst = torch.cuda.Stream()
        sst.wait_stream(torch.cuda.current_stream())
        with torch.cuda.graph(gt, stream=sst):
            tss = sharded.compress_slab(tiny, 64, 64, BZ * world, BX, BY, BZ, SCALE, LOCAL, out=tss)
        torch.cuda.current_stream().wait_stream(sst)
        barrier()
        t_fix = time_kernel(gt.replay, 50)
        barrier()
        t_fix_eager = time_kernel(lambda: sharded.compress_slab(tiny, 64, 64, BZ * world, BX, BY, BZ, SCALE, LOCAL, out=tss), 50)
        # the per-kernel times above are THIS rank's (rank 0 may own fewer block rows than the slowest rank): take the max over
        # ranks of both sums, as for the whole calls
        km = torch.tensor([sum(v["ms"] for k, v in kern.items() if not k.startswith("k_decode")), kern[dec_name]["ms"]],
                          dtype=torch.float64, device=dev)
        dist.all_reduce(km, op=dist.ReduceOp.MAX)
        ksum, kdec = float(km[0].item()), float(km[1].item())
        fixed = {"sharded_compress_of_a_64x64x%d_volume_ms" % (BZ * world): t_fix, "same_without_cuda_graph_ms": t_fix_eager,
                 "compress_ms": tc_ms, "compress_kernels_ms": ksum, "compress_overhead_ms": tc_ms - ksum,
                 "decompress_ms": td_ms, "decompress_kernel_ms": kdec,
                 "exchange": "peer-memory mailboxes (CUDA IPC over NVLink)" if L.cvxgpu_shard_uses_peer_memory(sharded.NcclShard.get().handle) else "2 ncclAllGather",
                 "note": "overhead = whole sharded call minus its two hot kernels timed alone (all three: max over ranks); the tiny-volume call is the floor of that overhead "
                         "(launch gaps + the exchange of 8..%d bytes per rank + the small kernels)" % (8 * max(1, -(-NZ // BZ // world)))}
        del gt

    # -- a short weak-scaling run beside the strong one: every rank a full volume's worth of planes --
    weak = None
    if world > 1 and args.scaling == "strong":
        del rt.graphs
        rt.graphs = None
        slab = out_slab = None
        del rt
        torch.cuda.empty_cache()
        rtw = RoundTrip(args, rank, world, dev, NZ * world, use_graph=not args.no_graph)
        wms, wtc, wtd, _ = rtw.timed(max(3, min(K, 5)), 2, barrier)
        weak = {"nz": NZ * world, "value": rtw.cells_total / (wms * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": wms,
                "compress_ms": wtc, "decompress_ms": wtd, "note": "one %d-plane slab per rank of a %d-plane volume" % (NZ, NZ * world)}
        del rtw

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        def ours_on_sample(vol_np):
            v = torch.from_numpy(vol_np).to(dev)
            s_, ln_, _ = api.compress_device(v, BX, BY, BZ, SCALE, LOCAL)
            r_ = api.decompress_device(s_, ln_, NX, NY, vol_np.shape[0])
            e_ = float(torch.sqrt(torch.sum((v.double() - r_.double()) ** 2) / torch.sum(v.double() ** 2)).item())
            return ln_, (-20.0 * math.log10(e_) if e_ > 0 else float("inf"))
        try:
            cpu, _, _ = cpu_leg(args.data, NZ, 3, 1, ours_on_sample)
        except Exception as e:       # the checker is optional at run time; say why it is missing
            cpu = {"value": None, "unit": UNIT, "cores": 0, "kind": "unavailable", "sample": repr(e)}
        gs = cpu.get("gpu_on_same_sample") if isinstance(cpu, dict) else None
        if gs and (abs(gs["snr_diff_db"]) > 0.1 or abs(gs["ratio_rel_diff"]) > 1e-3):
            sys.stderr.write("bench.py: INVALID RUN: on the reference arm's sample this path gives SNR %+.3f dB and ratio %+.4f %% "
                             "relative to the reference (bounds 0.1 dB, 0.1 %%)\n" % (gs["snr_diff_db"], 100 * gs["ratio_rel_diff"]))
            sys.exit(3)

    if rank == 0:
        # our kernels per step: [k_sumsq_rows, k_sumsq_rows_finish] (global RMS), k_rms_header, k_encode_prep, the encoder,
        # k_finish_length | k_shard_fixup (Compress); [k_shard_extent], the decoder (Decompress).  NCCL kernels are not counted.
        names = ([] if LOCAL else ["k_sumsq_rows", "k_sumsq_rows_finish"]) + ["k_rms_header", "k_encode_prep", enc_name,
                 "k_shard_fixup" if world > 1 else "k_finish_length"] + (["k_shard_extent"] if world > 1 else []) + [dec_name]
        line = {"metric": metric_name(args, world), "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": args.scaling if world > 1 else "strong", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "config": workload_config(args, world),
                "compress_mcells_s": cells_total / (tc_ms * 1e-3) / 1e6, "decompress_mcells_s": cells_total / (td_ms * 1e-3) / 1e6,
                "compress_ms": tc_ms, "decompress_ms": td_ms, "host_wall_ms_per_step": host_ms,
                "compressed_length": length, "ratio": 4.0 * cells_total / length, "roundtrip_error": err, "snr_db": snr,
                "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "e2e_pageable": e2e_pageable, "clocks": clocks,
                "gpu_launches": len(names) * K,
                "gpu_launches_detail": "per step and rank: " + ", ".join(names) + (" (+ 2 ncclAllGather)" if world > 1 else "") + (
                    "; each direction replayed from one CUDA graph" if (world > 1 and not args.no_graph) else ""),
                "exact_mode": exact_info, "weak": weak, "strong_scaling_fixed_cost": fixed, "numa": numa}
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def e2e_legs(args, rt, L, barrier, K):
    """The metric through the reference-facing calls on HOST buffers: `pinned` (the contract's e2e) and `pageable`
    (posix_memalign memory, what the reference's own drivers hand over: Test_With_Generated_Input.cpp:42-53)."""
    import numpy as np
    import torch
    world, dev, g = rt.world, rt.dev, rt.g
    slab, out_slab = rt.slab, rt.out_slab
    cells_rank, cells_total, length, nz_total = rt.cells_rank, rt.cells_total, rt.length, rt.nz_total
    ke = max(2, min(K, 5))
    res = []
    for kind in ("pinned", "pageable"):
        if kind == "pinned":
            h_vol = torch.empty(slab.shape, dtype=torch.float32, pin_memory=True)
            h_out = torch.empty(slab.shape, dtype=torch.float32, pin_memory=True)
        else:
            h_vol = torch.from_numpy(np.empty(slab.shape, dtype=np.float32))
            h_out = torch.from_numpy(np.zeros(slab.shape, dtype=np.float32))
        h_vol.copy_(slab)
        torch.cuda.synchronize()
        if world == 1:
            # exactly what a program linked against the reference calls (CvxCompress.hxx:169-192,
            # Test_With_Generated_Input.cpp:73-90): cvx_compress_th / cvx_decompress_inplace_th on host pointers
            h_stream = torch.empty(length + 64, dtype=torch.uint8, pin_memory=(kind == "pinned"))
            ln = C.c_long(0)

            def e2e_step():
                L.cvx_compress_th(SCALE, h_vol.data_ptr(), NX, NY, nz_total, BX, BY, BZ, bool(LOCAL), h_stream.data_ptr(), 1, C.byref(ln))
                L.cvx_decompress_inplace_th(h_out.data_ptr(), NX, NY, nz_total, h_stream.data_ptr(), 1, ln.value)
                return ln.value
            h2d = 4 * cells_rank + (length - 7)
            d2h = (length - 7) + 4 * cells_rank
            apiname = "cvx_compress_th + cvx_decompress_inplace_th (host pointers)"
        else:
            # a rank's slab comes from and returns to the host; the compressed shard stays on the GPU between the two
            # directions (a checkpoint that lives in HBM) -- nothing is bounced through the host for its own sake
            def e2e_step():
                slab.copy_(h_vol, non_blocking=True)                                   # H2D this rank's slab
                rt.compress()
                rt.decompress()
                h_out.copy_(out_slab, non_blocking=True)                               # D2H the decoded slab
                torch.cuda.synchronize()
                return 0
            h2d = 4 * cells_rank
            d2h = 4 * cells_rank
            apiname = "H2D slab + cvxgpu_compress_sharded / cvxgpu_decompress_sharded + D2H slab, per rank"
        e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(ke):
            e2e_step()
        torch.cuda.synchronize()
        t_e2e = (time.perf_counter() - t0) / ke
        tt = torch.tensor([t_e2e], dtype=torch.float64, device=dev)
        if world > 1:
            import torch.distributed as dist
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t_e2e = float(tt.item())
        same = bool(torch.equal(h_out, out_slab.cpu()))
        res.append({"value": cells_total / t_e2e / 1e6, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "ms_per_step": t_e2e * 1e3, "steps": ke, "host_memory": kind, "api": apiname,
                    "pcie_gbs_per_rank": (h2d + d2h) / t_e2e / 1e9, "matches_device_path": same})
        del h_vol, h_out
    return res[0], res[1]


# ---- BASELINE config 5: decompress-only, stream written by the reference on the host cores ------------------
def run_config5(args):
    """1280^3 G2 volume.  Rank 0 builds it on the host and lets the UNMODIFIED reference write the stream with all host
    cores (permuted block order, CvxCompress.cpp:365-390); every rank receives the whole stream and decodes its own iBlk
    range into its z-slab (cvxgpu_decompress_slab_device, no collective).  Checked against the reference's own decode of
    the same stream; the reverse direction (GPU stream -> reference decoder) is checked too."""
    import numpy as np
    import torch
    rank, world, local_rank, dev, numa = init_dist()
    import torch.distributed as dist
    from cvxcompress_b200 import _lib, api, sharded, synth
    L = _lib.lib()
    api.set_transform(args.transform)
    K, W = args.steps, args.warmup
    g = sharded.Geometry(NX, NY, NZ, BX, BY, BZ, rank, world)
    cells_total = NX * NY * NZ

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    cpu_info = None
    meta = torch.zeros(2, dtype=torch.int64, device=dev)
    host_vol = ref_out = None
    if rank == 0:
        host_vol = np.empty((NZ, NY, NX), np.float32)
        for z0 in range(0, NZ, 64):
            n = min(64, NZ - z0)
            host_vol[z0:z0 + n] = synth.g2_torch(NX, NY, n, dev, z0=z0, nz_total=NZ).cpu().numpy()
        cpu = CpuPath()
        cpu.prepare(host_vol)
        t_c = cpu.compress()
        t_d = cpu.decompress()
        ref_out = cpu.out.reshape(NZ, NY, NX)
        length = int(cpu.length.value)
        e_ref, snr_ref = quality_np(host_vol, ref_out)
        cpu_info = {"value": cells_total / t_d / 1e6, "unit": UNIT, "cores": cpu.cores, "kind": cpu.kind,
                    "compress_mcells_s": cells_total / t_c / 1e6, "decompress_mcells_s": cells_total / t_d / 1e6,
                    "sample": "the whole %dx%dx%d volume, one Compress (%.2f s) and one Decompress (%.2f s), threads=%d" % (
                        NX, NY, NZ, t_c, t_d, cpu.cores),
                    "compressed_length": length, "ratio": 4.0 * cells_total / length, "snr_db": snr_ref}
        meta[0] = length
        stream = torch.zeros((length + 15) // 16 * 16, dtype=torch.uint8, device=dev)
        stream[:length] = torch.from_numpy(np.ascontiguousarray(cpu.stream[:length])).to(dev)
    if world > 1:
        dist.broadcast(meta, src=0)
    length = int(meta[0].item())
    if rank != 0:
        stream = torch.zeros((length + 15) // 16 * 16, dtype=torch.uint8, device=dev)
    if world > 1:
        dist.broadcast(stream, src=0)
    out_slab = torch.empty((g.nz_slab, NY, NX), dtype=torch.float32, device=dev)
    hdr = (NX, NY, NZ, BX, BY, BZ, LOCAL)

    def do_decode():
        sharded.decompress_slab_from_stream(stream, length, rank=rank, world=world, out=out_slab, header=hdr)

    do_decode(); torch.cuda.synchronize()
    for _ in range(W):
        do_decode()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
    sampler = ClockSampler(local_rank)
    barrier(); sampler.start()
    ev[0].record()
    for k in range(K):
        do_decode(); ev[k + 1].record()
    torch.cuda.synchronize()
    clocks = sampler.stop(); barrier()
    tm = torch.tensor([ev[0].elapsed_time(ev[K]) / K], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    ms = float(tm.item())
    peak, peak_src = measured_peak_gbs()
    # ---- parity: every rank's slab against the reference's decode of the same stream ----
    ref_slab = torch.empty_like(out_slab)
    if world > 1:
        if rank == 0:
            for r in range(world - 1, -1, -1):
                gr = sharded.Geometry(NX, NY, NZ, BX, BY, BZ, r, world)
                t = torch.from_numpy(ref_out[gr.z0:gr.z1]).to(dev)
                if r == 0:
                    ref_slab.copy_(t)
                else:
                    dist.send(t, dst=r)
        else:
            dist.recv(ref_slab, src=0)
    else:
        ref_slab.copy_(torch.from_numpy(ref_out))
    st = torch.stack([(out_slab - ref_slab).abs().max().double(), ref_slab.abs().max().double(),
                      (out_slab != ref_slab).sum().double()])
    if world > 1:
        mx = st[:2].clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = st[2:].clone(); dist.all_reduce(sm)
        st = torch.cat([mx, sm])
    maxdiff, maxref, ndiff = (float(v) for v in st.tolist())
    tol = 1e-6 if args.transform == "exact" else 2e-5
    if not (maxdiff <= tol * maxref):
        sys.stderr.write("bench.py: INVALID RUN: GPU decode of the reference stream differs from the reference's decode by %g (max |v| %g)\n" % (maxdiff, maxref))
        sys.exit(3)
    # ---- reverse direction: a GPU-written stream through the reference decoder (rank 0, N == 1 only: it needs the volume on one GPU) ----
    reverse = None
    if world == 1:
        vol_d = torch.from_numpy(host_vol).to(dev)
        s2, ln2, _ = api.compress_device(vol_d, BX, BY, BZ, SCALE, LOCAL)
        gpu_dec = api.decompress_device(s2, ln2, NX, NY, NZ)
        h_s2 = s2[:ln2].cpu().numpy()
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_libs as ol
        if ol.have_ref():
            buf = ol.aligned_u8(ln2 + 64); buf[:ln2] = h_s2
            out2 = np.zeros(cells_total, np.float32)
            ol.ref().cvx_decompress_inplace_th(out2, NX, NY, NZ, buf, host_cores(), ln2)
            d2 = float((gpu_dec - torch.from_numpy(out2.reshape(NZ, NY, NX)).to(dev)).abs().max().item())
            reverse = {"gpu_stream_bytes": int(ln2), "max_abs_diff_reference_decode_vs_gpu_decode": d2, "max_abs_value": maxref,
                       "within": tol * maxref, "ok": d2 <= tol * maxref,
                       "length_rel_diff_vs_reference_stream": (length - ln2) / float(length)}
            if not reverse["ok"]:
                sys.stderr.write("bench.py: INVALID RUN: reference decode of the GPU stream differs from the GPU decode by %g\n" % d2)
                sys.exit(3)
        del vol_d, s2, gpu_dec
    # ---- e2e: host stream in, host slab out ----
    e2e = None
    if not args.no_e2e:
        h_stream = torch.empty(length + 64, dtype=torch.uint8, pin_memory=True)
        h_stream[:length].copy_(stream[:length])
        h_out = torch.empty(out_slab.shape, dtype=torch.float32, pin_memory=True)
        torch.cuda.synchronize()
        if world == 1:
            def step():
                L.cvx_decompress_inplace_th(h_out.data_ptr(), NX, NY, NZ, h_stream.data_ptr(), 1, length)
        else:
            def step():
                stream[:length].copy_(h_stream[:length], non_blocking=True)
                do_decode()
                h_out.copy_(out_slab, non_blocking=True)
                torch.cuda.synchronize()
        step(); barrier()
        ke = max(2, min(K, 4))
        t0 = time.perf_counter()
        for _ in range(ke):
            step()
        torch.cuda.synchronize()
        tt = torch.tensor([(time.perf_counter() - t0) / ke], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        te = float(tt.item())
        e2e = {"value": cells_total / te / 1e6, "unit": UNIT, "h2d_bytes_per_step": int(length), "d2h_bytes_per_step": int(4 * out_slab.numel()),
               "ms_per_step": te * 1e3, "steps": ke, "host_memory": "pinned",
               "api": "cvx_decompress_inplace_th (host pointers)" if world == 1 else "H2D stream + cvxgpu_decompress_slab_device + D2H slab, per rank"}
    if rank == 0:
        tb = 8 * g.nblk
        abytes = 4 * out_slab.numel() + (length - 32 - tb) * (g.nblk_slab / float(g.nblk)) + 8 * g.nblk_slab
        gbs = abytes / (ms * 1e-3) / 1e9
        line = {"metric": metric_name(args, world), "value": cells_total / (ms * 1e-3) / 1e6, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": workload_config(args, world), "compressed_length": length, "ratio": 4.0 * cells_total / length,
                "parity": {"max_abs_diff_vs_reference_decode": maxdiff, "max_abs_value": maxref, "cells_differing": ndiff,
                           "tolerance": tol * maxref, "stream_block_order": "reference, %d threads (permuted)" % (cpu_info or {}).get("cores", 0),
                           "reverse": reverse},
                "roofline": {"bound": "hbm", "kernel": "k_decode32ws" if args.transform == "lifting" else "k_decode32", "achieved": gbs, "peak": peak,
                             "unit": "GB/s", "frac": gbs / peak, "traffic": None, "peak_source": peak_src,
                             "algorithmic_bytes_per_launch": abytes, "ms_per_launch": ms},
                "cpu_baseline": cpu_info, "e2e": e2e, "clocks": clocks, "gpu_launches": K,
                "gpu_launches_detail": "per step and rank: one decode kernel over the rank's iBlk range", "numa": numa}
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# ---- BASELINE config 2: wavelet forward + inverse sweep over cubes 8^3 .. 256^3 ------------------------------
SWEEP = (8, 16, 32, 64, 128, 256)


def run_wavelet(args):
    """CvxCompress_Test's throughput sweep (CvxCompress.cpp:803-891): forward then inverse 9/7 transform of dense cubes.
    The reference re-transforms ONE cache-resident block per thread (2^30 cells' worth of repetitions); here 2^30 cells of
    DISTINCT blocks (4 GiB) stream through HBM: 16*N algorithmic bytes per forward + inverse pair."""
    import torch
    rank, world, local_rank, dev, numa = init_dist()
    from cvxcompress_b200 import api, synth
    api.set_transform(args.transform)
    K, W = args.steps, args.warmup
    ncell = 1 << 30
    if args.nz:
        ncell = max(1 << 24, (args.nz * 1056 * 1248) // (1 << 24) * (1 << 24))
    peak, peak_src = measured_peak_gbs()
    idx = torch.arange(ncell, dtype=torch.int64, device=dev)
    src = ((synth._hash32_torch(idx) >> 8).to(torch.float32) / float(1 << 24)) * 2.0 - 1.0       # U(-1, 1), as Fill_Block's range
    del idx
    buf = torch.empty_like(src)
    sweep = {}
    sampler = ClockSampler(local_rank)
    sampler.start()
    for b in SWEEP:
        nblocks = ncell // (b * b * b)
        blocks = buf.view(nblocks, b, b, b)
        buf.copy_(src)
        api.wavelet_blocks_device(blocks, b, b, b, inverse=False)
        api.wavelet_blocks_device(blocks, b, b, b, inverse=True)
        torch.cuda.synchronize()
        err = float((buf - src).abs().max().item())
        tf = time_kernel(lambda: api.wavelet_blocks_device(blocks, b, b, b, inverse=False), max(2, min(K, 5)))
        ti = time_kernel(lambda: api.wavelet_blocks_device(blocks, b, b, b, inverse=True), max(2, min(K, 5)))
        gbs = 16.0 * ncell / ((tf + ti) * 1e-3) / 1e9
        sweep["%d^3" % b] = {"forward_ms": tf, "inverse_ms": ti, "mcells_s": ncell / ((tf + ti) * 1e-3) / 1e6,
                             "achieved_gbs": gbs, "frac": gbs / peak, "max_roundtrip_error": err, "blocks": nblocks}
    clocks = sampler.stop()
    cpu = None
    if rank == 0 and not args.no_cpu:
        try:
            cpu = wavelet_cpu_leg()
        except Exception as e:
            cpu = {"value": None, "unit": UNIT, "cores": 0, "kind": "unavailable", "sample": repr(e)}
    if rank == 0:
        top = sweep["32^3"]
        line = {"metric": "Wavelet forward+inverse MCells/s, cubes 8^3..256^3, %d cells of distinct blocks (value = 32^3)" % ncell,
                "value": top["mcells_s"], "unit": UNIT, "n_gpus": 1, "steps": K, "warmup": W, "ms_per_step": top["forward_ms"] + top["inverse_ms"],
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": "BASELINE config 2: CvxCompress_Test wavelet fwd+inv throughput sweep, block sizes 8^3..256^3, "
                                       "G3 data U(-1,1), %d cells (%.1f GiB) of distinct dense blocks per size" % (ncell, ncell * 4 / 2 ** 30),
                           "config": 2, "transform": args.transform, "cells": ncell,
                           "l2": "inputs larger than L2 (4 GiB per pass vs 126 MB L2); no flush needed"},
                "sweep": sweep,
                "roofline": {"bound": "hbm", "kernel": "k_wavelet_fused (32^3)", "achieved": top["achieved_gbs"], "peak": peak, "unit": "GB/s",
                             "frac": top["frac"], "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_launch": 8 * ncell,
                             "ms_per_launch": 0.5 * (top["forward_ms"] + top["inverse_ms"])},
                "cpu_baseline": cpu, "e2e": None, "clocks": clocks, "gpu_launches": 2 * len(SWEEP) * max(2, min(K, 5)), "numa": numa}
        emit(line)


def wavelet_cpu_leg(budget_s=3.0):
    """The reference's Wavelet_Transform_Fast_{Forward,Inverse} (oracle/_ref) on one cube per size, one thread, repeated for
    ~budget_s / size -- its own benchmark's cache-resident regime (CvxCompress.cpp:826-857)."""
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_libs as ol
    if not ol.have_ref():
        raise RuntimeError("oracle/_ref not built")
    R = ol.ref()
    from cvxcompress_b200 import synth
    res = {}
    for b in SWEEP:
        n = b * b * b
        blk = ol.aligned_f32(n)
        blk[:] = synth.g3(n, lo=-1.0, hi=1.0)
        work = ol.aligned_f32(max(n, 1 << 16))
        reps, t0 = 0, time.perf_counter()
        while time.perf_counter() - t0 < budget_s / len(SWEEP) or reps < 2:
            R.fwd_fast(blk.ctypes.data, work.ctypes.data, b, b, b)
            R.inv_fast(blk.ctypes.data, work.ctypes.data, b, b, b)
            reps += 1
        dt = time.perf_counter() - t0
        res["%d^3" % b] = {"mcells_s": reps * n / dt / 1e6, "pairs": reps}
    return {"value": res["32^3"]["mcells_s"], "unit": UNIT, "cores": 1, "kind": "reference",
            "sample": "Wavelet_Transform_Fast_Forward + _Inverse of one cache-resident cube per size, 1 thread, ~%.1f s per size" % (budget_s / len(SWEEP)),
            "sweep": res}


def run_reference_wavelet(args):
    info = wavelet_cpu_leg(budget_s=max(6.0, 2.0 * args.steps))
    emit({"impl": "reference", "metric": "Wavelet forward+inverse MCells/s, cubes 8^3..256^3 (value = 32^3)", "value": info["value"], "unit": UNIT,
          "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": None, "higher_is_better": True, "scaling": "strong",
          "vs_baseline": None, "dtype": "f32", "data": "synthetic",
          "config": {"workload": "BASELINE config 2 on the host: " + info["sample"], "config": 2}, "cpu_baseline": info,
          "e2e": {"value": info["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0})


_RESULT_FD = None


def _claim_stdout():
    """stdout carries exactly ONE JSON line (rank 0).  Libraries write there too (NCCL prints its "NCCL version ..." banner
    on fd 1 when the box sets NCCL_DEBUG), so fd 1 is pointed at stderr for the life of the process and the result line
    goes to a private duplicate of the original stdout."""
    global _RESULT_FD
    if _RESULT_FD is None:
        sys.stdout.flush()
        _RESULT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def main():
    args = parse_args()
    select_config(args)
    _claim_stdout()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
