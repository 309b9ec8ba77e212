#!/usr/bin/env python
"""bench.py -- headline benchmark of the b200jpeg decode path (BASELINE.json: decoded Mpixel/s & images/s, batched JPEG
1080p 4:2:0).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--batch 64]

One step = one pass of the hot path over one batch: 64 synthetic 1920x1080 4:2:0 q90 baseline JPEGs (no restart
markers) -> interleaved RGB uint8 (BASELINE.json configs[0]/[1] shape: the configuration the metric is quoted on).
  value : whole-job Mpixel/s with the compressed bytes already resident in HBM (b200jpeg_decode_run only: every kernel)
  e2e   : the same metric through the C ABI call a plugin user makes, HOST jpeg bytes in -> HOST pixels out
          (b200jpeg_decode_batch with on_host outputs: parse, pinned staging, H2D, kernels, D2H inside the timed region)
  roofline     : dominant kernel stage, algorithmic bytes / CUDA-event time vs the measured HBM peak
  cpu_baseline : the reference's own CPU path (libjpeg_turbo_ext through nvimgcodecDecoderDecode, oracle/_ref build)
N>1: one process per GPU (torchrun), the batch is replicated per rank (weak scaling), no collective on the data path.
"""
import argparse
import io
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H, QUALITY = 1920, 1080, 90


def _encode_one(seed):
    from PIL import Image
    from nvimagecodec_b200.corpus import synth_image
    b = io.BytesIO()
    Image.fromarray(synth_image(seed, W, H)).save(b, "JPEG", quality=QUALITY, subsampling="4:2:0", optimize=False)
    return b.getvalue()


def make_corpus(n):
    """n seeded synthetic 1080p JPEGs; encoded by PIL's bundled libjpeg-turbo (byte-identical to the reference's
    libjpeg-turbo 3.0.3 jpeg_compress with the same parameters -- checked in tests/test_oracle.py)"""
    from concurrent.futures import ProcessPoolExecutor
    import multiprocessing as mp
    workers = max(1, min(16, (os.cpu_count() or 1)))
    with ProcessPoolExecutor(max_workers=workers, mp_context=mp.get_context("spawn")) as ex:
        return list(ex.map(_encode_one, [1000 + i for i in range(n)]))


class ClockSampler:
    """SM clocks and throttle reasons of the job's GPUs during the timed region.  One sampler for the whole job (rank 0
    passes every GPU index, the other ranks none): NVML in-process every 10 ms; eight ranks each spawning nvidia-smi --
    which opens every GPU of the box -- inside a 30 ms timed region is a disturbance of its own.  Falls back to one
    nvidia-smi process per sample (all GPUs in one query) when pynvml is missing."""
    FIELDS = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    BITS = [0x8, 0x40, 0x20, 0x4]  # nvmlClocksEventReason{HwSlowdown, HwThermalSlowdown, SwThermalSlowdown, SwPowerCap}

    def __init__(self, indices):
        self.indices = [indices] if isinstance(indices, int) else list(indices)
        self.samples = []  # (gpu, sm_mhz, max_mhz, [reason flags])
        self.stop = False
        self.how = None
        self.thread = threading.Thread(target=self._run, daemon=True) if self.indices else None

    def _run_nvml(self):
        import pynvml
        pynvml.nvmlInit()
        hs = [pynvml.nvmlDeviceGetHandleByIndex(i) for i in self.indices]
        mx = [pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM) for h in hs]
        reasons_of = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
        self.how = "nvml"
        while not self.stop:
            for g, h in enumerate(hs):
                sm = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
                r = int(reasons_of(h))
                self.samples.append((self.indices[g], int(sm), int(mx[g]), [bool(r & b) for b in self.BITS]))
            time.sleep(0.01)

    def _run_smi(self):
        self.how = "nvidia-smi"
        while not self.stop:
            try:
                out = subprocess.run(["nvidia-smi", "-i", ",".join(str(i) for i in self.indices), "--query-gpu=" + self.FIELDS,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                for g, line in enumerate(out.strip().splitlines()):
                    parts = [p.strip() for p in line.split(",")]
                    if len(parts) >= 6 and parts[0].isdigit() and parts[1].isdigit() and g < len(self.indices):
                        self.samples.append((self.indices[g], int(parts[0]), int(parts[1]),
                                             [parts[2 + i].lower().startswith("active") for i in range(4)]))
            except Exception:
                pass
            time.sleep(0.1)

    def _run(self):
        try:
            self._run_nvml()
        except Exception:
            self._run_smi()

    def __enter__(self):
        if self.thread:
            self.thread.start()
            time.sleep(0.05)  # the first sample is in before the timed region starts
        return self

    def __exit__(self, *a):
        self.stop = True
        if self.thread:
            self.thread.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        per_gpu = {}
        for g, sm, _, _ in self.samples:
            per_gpu.setdefault(g, []).append(sm)
        med = {g: sorted(v)[len(v) // 2] for g, v in per_gpu.items()}
        reasons = [n for i, n in enumerate(self.NAMES) if any(s[3][i] for s in self.samples)]
        out = {"sm_mhz": min(med.values()), "sm_max_mhz": max(s[2] for s in self.samples), "reasons": reasons,
               "samples": len(self.samples), "how": self.how}
        if len(med) > 1:
            out["per_gpu_sm_mhz"] = [med[g] for g in sorted(med)]  # sm_mhz is the lowest of these medians
        return out


def _gpu_local_cpus(pynvml, index):
    try:
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(index)).busId
        bus = (bus.decode() if isinstance(bus, bytes) else bus).lower()
        if len(bus.split(":")[0]) == 8:
            bus = bus[4:]
        return open("/sys/bus/pci/devices/" + bus + "/local_cpulist").read().strip()
    except Exception:
        return None


def pin_to_gpu_numa_node(index):
    """one host thread set per device: run this rank on the cores next to its GPU so that pinned staging
    buffers are allocated on, and copied from, the local memory node.  Returns a description or None."""
    try:
        import pynvml
        pynvml.nvmlInit()
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(index)).busId
        bus = (bus.decode() if isinstance(bus, bytes) else bus).lower()
        if len(bus.split(":")[0]) == 8:
            bus = bus[4:]
        base = "/sys/bus/pci/devices/" + bus
        node = int(open(base + "/numa_node").read())
        cpus = set()
        for part in open(base + "/local_cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            # the ranks whose GPUs share these cores take disjoint slices of them (the decoder sizes its staging threads by
            # the cores its process may use): without that, N ranks x (copy threads + completion threads) fight for the box
            world = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1")))
            sharing = [r for r in range(world) if _gpu_local_cpus(pynvml, r) == _gpu_local_cpus(pynvml, index)] or [index]
            ordered = sorted(cpus)
            per = max(1, len(ordered) // len(sharing))
            k = sharing.index(index) if index in sharing else 0
            mine = set(ordered[k * per:(k + 1) * per]) or cpus
            os.sched_setaffinity(0, mine)
            return {"numa_node": node, "cpus": len(mine), "shared_by_ranks": len(sharing)}
    except Exception:
        pass
    return None


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def workload_string(args):
    if args.workload == "c2":
        return "C2: %dx variable-size ~500x375 4:2:0 q75 baseline JPEG -> interleaved RGB u8, fancy upsampling" % args.batch
    return "C1: %dx 1920x1080 4:2:0 q90 baseline JPEG, no restart markers -> interleaved RGB u8, fancy upsampling" % args.batch


def workload_config(args):
    """identical in both arms (the driver compares them)"""
    return {"workload": workload_string(args), "batch": args.batch, "options": ":fancy_upsampling=1"}


def _encode_c2(spec):
    from PIL import Image
    from nvimagecodec_b200.corpus import synth_image
    seed, w, h = spec
    b = io.BytesIO()
    Image.fromarray(synth_image(seed, w, h)).save(b, "JPEG", quality=75, subsampling="4:2:0", optimize=False)
    return b.getvalue()


def make_workload(args):
    """-> (list of JPEG bytes, list of (w, h)); C2 repeats 64 distinct images up to the batch size"""
    if args.workload != "c2":
        return make_corpus(args.batch), [(W, H)] * args.batch
    from concurrent.futures import ProcessPoolExecutor
    import multiprocessing as mp
    rng = np.random.default_rng(2)
    distinct = min(args.batch, 64)
    specs = [(2000 + i, int(rng.integers(320, 641)), int(rng.integers(240, 501))) for i in range(distinct)]
    with ProcessPoolExecutor(max_workers=max(1, min(16, os.cpu_count() or 1)), mp_context=mp.get_context("spawn")) as ex:
        base = list(ex.map(_encode_c2, specs))
    return [base[i % distinct] for i in range(args.batch)], [(specs[i % distinct][1], specs[i % distinct][2]) for i in range(args.batch)]


def reference_decode_all(datas, dims, reps, threads=0):
    """reference CPU path: nvimgcodecDecoderDecode -> libjpeg_turbo_ext on the thread-pool executor, host RGB out"""
    from oracle import ref
    sess = ref.Session.shared()
    dec = sess.decoder(device_id=ref.DEVICE_CPU_ONLY, backends=ref.BACKEND_CPU_ONLY, num_threads=threads, options=":fancy_upsampling=1")
    outs, st, secs = dec.decode_host(datas, fmt=ref.I_RGB, reps=reps, dims=dims)
    assert all(s == ref.STATUS_SUCCESS for s in st)
    return outs, secs


def simd_context(datas, dims, cores):
    """What WITH_SIMD=0 costs the reference arm: the same bytes through a stock SIMD build of libjpeg-turbo (the one PIL bundles;
    islow IDCT and fancy upsampling are its defaults too) on every host thread.  Context only -- not the reference's code path."""
    try:
        import PIL
        from PIL import Image
        from concurrent.futures import ThreadPoolExecutor

        def one(d):
            im = Image.open(io.BytesIO(d))
            im.load()  # releases the GIL while libjpeg runs
            return im.size

        with ThreadPoolExecutor(max_workers=cores) as tp:
            list(tp.map(one, datas[:cores]))
            best = None
            for _ in range(3):
                t0 = time.perf_counter()
                list(tp.map(one, datas))
                dt = time.perf_counter() - t0
                best = dt if best is None else min(best, dt)
        feats = ""
        try:
            from PIL import features
            feats = " libjpeg-turbo %s" % features.version_feature("libjpeg_turbo") if features.check_feature("libjpeg_turbo") else ""
        except Exception:
            pass
        return {"value": sum(w * h for w, h in dims) / 1e6 / best, "unit": "Mpixel/s", "cores": cores,
                "how": "PIL %s%s (stock SIMD build), Image.load() of the same %d files on %d threads, best of 3; context only" %
                       (PIL.__version__, feats, len(datas), cores)}
    except Exception as e:  # no PIL on the box: say so
        return {"value": None, "how": "unavailable: %s" % e}


def run_reference(args, rank, world):
    if rank != 0:
        return
    n = args.batch
    datas, dims = make_workload(args)
    cores = os.cpu_count() or 1
    mpix = sum(w * h for w, h in dims) / 1e6
    _, warm = reference_decode_all(datas, dims, max(1, min(args.warmup, 2)))
    _, secs = reference_decode_all(datas, dims, args.steps)
    total = float(sum(secs))
    value = mpix * args.steps / total
    line = {
        "impl": "reference", "metric": "decoded_mpixel_per_s", "value": value, "unit": "Mpixel/s", "images_per_s": n * args.steps / total,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True,
        "scaling": "strong" if args.workload == "c2" else "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(args),
        "reference_path": "nvimgcodecDecoderDecode -> libjpeg_turbo_ext (libjpeg-turbo 3.0.3, WITH_SIMD=0: no nasm in the image) on the "
                          "default thread-pool executor, host buffers; rank 0 only, every image of the batch in every step",
        "cpu_baseline": {"value": value, "unit": "Mpixel/s", "cores": cores, "kind": "reference",
                         "sample": "all %d images per step, %d steps, all %d host threads" % (n, args.steps, cores),
                         "simd_context": simd_context(datas, dims, cores)},
        "e2e": {"value": value, "unit": "Mpixel/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def _synth_one(seed):
    from nvimagecodec_b200.corpus import synth_image
    return synth_image(seed, W, H)


def run_encode(args, rank, world, local_rank):
    """BASELINE.json configs[4]: batched JPEG encode, 128x 1920x1080 RGB -> 4:2:0 q90 baseline, bitstream-exact vs
    libjpeg-turbo.  value: device-resident pixels -> host JPEG bytes; e2e: pinned host pixels -> host JPEG bytes."""
    from concurrent.futures import ProcessPoolExecutor, ThreadPoolExecutor
    import multiprocessing as mp
    n = args.batch
    distinct = min(n, 16)
    with ProcessPoolExecutor(max_workers=max(1, min(16, os.cpu_count() or 1)), mp_context=mp.get_context("spawn")) as ex:
        base = list(ex.map(_synth_one, [2000 + i for i in range(distinct)]))
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the b200jpeg path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ["NCCL_DEBUG"] = "WARN"
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
            torch.cuda.synchronize(dev)
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    from nvimagecodec_b200 import _capi
    from nvimagecodec_b200.encoder import Encoder
    from oracle import port, ref
    host = [torch.from_numpy(base[i % distinct]).pin_memory() for i in range(n)]
    devt = [h.to(dev) for h in host]
    enc = Encoder(local_rank)
    encp = Encoder(local_rank, options=":profile=1")  # single stream, CUDA events between the stages
    cap = W * H * 3 // 2 + 4096
    caps = [cap] * n
    out_host = torch.empty(n * cap, dtype=torch.uint8).pin_memory()  # page-locked destination of the JPEG files
    out_ptrs = [out_host.data_ptr() + i * cap for i in range(n)]

    def inputs(tensors, on_host):
        return [dict(buffer=t.data_ptr(), row_stride=W * 3, width=W, height=H, format=_capi.FMT_I_RGB, on_host=int(on_host),
                     subsampling=_capi.CSS_420, quality=QUALITY, restart_interval=0) for t in tensors]

    # ---- parity gate + CPU baseline: libjpeg-turbo jpeg_compress (the reference tree's encoder) on the host cores ----
    torch.cuda.synchronize(dev)
    got, st = enc.encode_raw(inputs(devt, False), caps)
    assert all(x == 0 for x in st), st
    cpu_baseline = None
    if rank == 0:
        lj = (lambda im: ref.ljt_encode(im, QUALITY, (2, 2))) if ref.available() else (lambda im: port.encode(im, QUALITY, (2, 2)))
        kind = "reference" if ref.available() else "port"
        cores = os.cpu_count() or 1
        for i in range(distinct):
            if got[i] != lj(base[i]):
                raise SystemExit("PARITY FAILURE: encoded image %d differs from libjpeg-turbo" % i)
        assert all(got[i] == got[i % distinct] for i in range(n))
        if not args.skip_cpu_baseline:
            sample = [base[i % distinct] for i in range(min(n, max(16, 2 * cores)))]
            with ThreadPoolExecutor(max_workers=cores) as tp:  # ctypes releases the GIL: one jpeg_compress per host thread
                list(tp.map(lj, sample[:cores]))
                t0 = time.perf_counter()
                list(tp.map(lj, sample))
                dt = time.perf_counter() - t0
            cpu_baseline = {"value": len(sample) * W * H / 1e6 / dt, "unit": "Mpixel/s", "cores": cores, "kind": kind,
                            "images_per_s": len(sample) / dt,
                            "sample": "%d of the %d images, libjpeg-turbo 3.0.3 jpeg_compress (WITH_SIMD=0) on %d host threads" % (len(sample), n, cores)}

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(inp, enc=enc):
        for _ in range(max(3, args.warmup)):
            enc.encode_into(inp, out_ptrs, caps)
        barrier()
        stage = [0.0, 0.0, 0.0]
        launches = 0
        t0 = time.perf_counter()
        for _ in range(args.steps):
            sizes, st = enc.encode_into(inp, out_ptrs, caps)
            es = enc.stats()
            launches += es["launches"]
            stage = [stage[0] + es["dct_us"], stage[1] + es["entropy_us"], stage[2] + es["stuff_us"]]
        torch.cuda.synchronize(dev)
        dt = time.perf_counter() - t0
        assert all(x == 0 for x in st)
        if rank == 0:  # the timed path wrote the same files as the parity-checked call
            for i in (0, n - 1):
                assert out_host[i * cap:i * cap + sizes[i]].numpy().tobytes() == got[i]
        return dt, [v / args.steps / 1e3 for v in stage], launches, enc.stats()

    with ClockSampler(range(world) if rank == 0 else []) as clocks:
        dt_dev, _, launches, es = timed(inputs(devt, False))
    dt_host, _, _, _ = timed(inputs(host, True))
    _, stage_ms, _, _ = timed(inputs(devt, False), encp)
    t = torch.tensor([dt_dev, dt_host], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dt_dev, dt_host = float(t[0]), float(t[1])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    mpix = n * W * H / 1e6 * world
    peak, peak_kind = hbm_peak()
    blocks, in_b, out_b = es["blocks"], es["in_bytes"], es["out_bytes"]
    # stage 0 = k_enc_color + k_enc_fdct: reads the pixels once, writes and re-reads the component planes (64 B per block), writes
    # 128 B of zigzag coefficients + 12 B (DC, AC length, non-zero mask) per block
    alg_dct = in_b + (2 * 64 + 140) * blocks
    line = {
        "metric": "encoded_mpixel_per_s", "value": mpix / (dt_dev / args.steps), "unit": "Mpixel/s", "images_per_s": n * world / (dt_dev / args.steps),
        "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": 1e3 * dt_dev / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "C5: %dx 1920x1080 interleaved RGB u8 -> 4:2:0 q90 baseline JPEG (standard Huffman tables, %.0f KB avg), per GPU"
                               % (n, out_b / n / 1e3),
                   "l2": "working set per step %.0f MB (pixels+coefficients) > 126 MB L2, no explicit flush" % ((in_b + 128 * blocks) / 1e6),
                   "parity": "byte-exact vs libjpeg-turbo jpeg_compress on all distinct images before timing",
                   "timing": "wall clock around b200jpeg_encode_batch (synchronous; per group of images it reads the stream lengths back "
                             "mid-way), device-resident pixels in, JPEG files in pinned host memory out"},
        "clocks": clocks.summary(),
        "e2e": {"value": mpix / (dt_host / args.steps), "unit": "Mpixel/s", "images_per_s": n * world / (dt_host / args.steps),
                "ms_per_step": 1e3 * dt_host / args.steps, "h2d_bytes_per_step": int(in_b), "d2h_bytes_per_step": int(out_b)},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": "k_enc_color + k_enc_fdct", "achieved": alg_dct / (stage_ms[0] * 1e-3) / 1e9 if stage_ms[0] else None, "peak": peak,
                     "peak_kind": peak_kind, "unit": "GB/s", "frac": alg_dct / (stage_ms[0] * 1e-3) / 1e9 / peak if stage_ms[0] else None,
                     "traffic": None, "algorithmic_bytes": alg_dct, "ms": stage_ms[0],
                     "stage_ms": {"color+fdct (incl. H2D of descriptors)": stage_ms[0], "bit offsets + emit": stage_ms[1], "stuff": stage_ms[2]}},
        "cpu_baseline": cpu_baseline,
    }
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--options", default=":fancy_upsampling=1")
    ap.add_argument("--skip-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="decode", choices=["decode", "c2", "encode"],
                    help="decode = the headline C1 decode batch (default); c2 = BASELINE configs[1], 256 ImageNet-shaped JPEGs dealt over the "
                         "GPUs (strong scaling); encode = BASELINE configs[4], 128x 1080p RGB -> JPEG")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.workload == "encode":
        if args.impl == "reference":
            raise SystemExit("the reference arm is defined for the decode headline; the encode line carries its libjpeg-turbo baseline")
        if args.batch == 64:
            args.batch = 128
        run_encode(args, rank, world, local_rank)
        return
    if args.workload == "c2" and args.batch == 64:
        args.batch = 256
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    run_decode(args, rank, world, local_rank)


def run_decode(args, rank, world, local_rank):
    datas_all, dims_all = make_workload(args)  # before CUDA init (spawned workers)
    numa = pin_to_gpu_numa_node(local_rank) if world > 1 else None
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the b200jpeg path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # NCCL announces its version on stdout at the first collective: keep stdout for the one JSON line
        os.environ["NCCL_DEBUG"] = "WARN"
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
            torch.cuda.synchronize(dev)
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    import nvimagecodec_b200 as nb
    from nvimagecodec_b200.decoder import Decoder
    from nvimagecodec_b200.sharding import estimate_cost, shard_batch
    from oracle import ref  # the unmodified reference FRAMEWORK is the caller of the plugin in the e2e legs (and the checker)

    # C1: every rank decodes the whole batch (weak scaling); C2: the batch is dealt over the ranks (strong scaling, fixed N)
    sharded = args.workload == "c2"
    if sharded and world > 1:
        mine = shard_batch([estimate_cost(len(d), w, h) for d, (w, h) in zip(datas_all, dims_all)], world)[rank]
    else:
        mine = list(range(len(datas_all)))
    datas = [datas_all[i] for i in mine]
    dims = [dims_all[i] for i in mine]
    n = len(datas)
    pixels = sum(w * h for w, h in dims)
    out_bytes = 3 * pixels
    opts = args.options
    stream = torch.cuda.current_stream(dev).cuda_stream

    # device outputs: the images of a set back to back in one allocation (a batch tensor), two sets for the two batches in flight
    offs = np.concatenate([[0], np.cumsum([3 * w * h for w, h in dims])]).astype(np.int64)
    dev_sets = [torch.empty(out_bytes + 64, dtype=torch.uint8, device=dev) for _ in range(2)]
    host_sets = [torch.empty(out_bytes + 64, dtype=torch.uint8).pin_memory() for _ in range(2)]
    ptrs = [[t.data_ptr() + int(offs[i]) for i in range(n)] for t in dev_sets]
    hptrs = [[t.data_ptr() + int(offs[i]) for i in range(n)] for t in host_sets]
    strides = [3 * w for w, h in dims]
    fmts = [4] * n

    dec = Decoder(local_rank, options=opts)
    st = dec.decode_into(datas, ptrs[0], strides, fmts, False, stream)
    assert all(s == 0 for s in st), st
    torch.cuda.synchronize(dev)

    # ---- parity gate + CPU baseline (rank 0): the same bytes through the reference's CPU path ----
    cpu_baseline = None
    want = None
    if rank == 0 and not args.skip_cpu_baseline:
        cores = len(os.sched_getaffinity(0)) or 1  # at N > 1 this rank owns a slice of the host's cores (pin_to_gpu_numa_node)
        reps = 3 if cores >= 16 else 1
        want, secs = reference_decode_all(datas, dims, reps)
        got = dev_sets[0].cpu().numpy()
        for i, (w, h) in enumerate(dims):
            if not np.array_equal(got[offs[i]:offs[i + 1]].reshape(h, w, 3), want[i]):
                raise SystemExit("PARITY FAILURE on image %d: GPU output differs from the reference decoder" % i)
        best = min(secs)
        cpu_baseline = {"value": pixels / 1e6 / best, "unit": "Mpixel/s", "cores": cores, "kind": "reference", "images_per_s": n / best,
                        "sample": "all %d images of this rank's batch, best of %d passes, nvimgcodec default executor with %d threads, "
                                  "libjpeg-turbo 3.0.3 WITH_SIMD=0" % (n, reps, cores),
                        "simd_context": simd_context(datas, dims, cores)}

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- value: device-resident inputs, kernels only ----
    st = dec.prepare(datas, ptrs[0], strides, fmts, False)
    assert all(s == 0 for s in st)
    for _ in range(args.warmup):
        dec.run(stream)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches = 0
    with ClockSampler(range(world) if rank == 0 else []) as clocks:
        e0.record()
        for _ in range(args.steps):
            st = dec.run(stream)
            launches += dec.stats()["launches"]
        e1.record()
        barrier()
    assert all(s == 0 for s in st)
    ms_total = e0.elapsed_time(e1)
    stats = dec.stats()

    # ---- per-stage device times: same batch, same steps, single stream with CUDA events between the stages ----
    stage_acc = {}
    decp = Decoder(local_rank, options=opts + " :profile=1")
    st = decp.prepare(datas, ptrs[0], strides, fmts, False)
    for _ in range(args.warmup):
        decp.run(stream)
    for _ in range(args.steps):
        decp.run(stream)
        for k, v in (decp.stage_times() or {}).items():
            stage_acc[k] = stage_acc.get(k, 0.0) + v
    decp.close()
    barrier()

    # ---- C ABI, host JPEG bytes -> host RGB (b200jpeg_decode_batch) ----
    for _ in range(2):
        dec.decode_into(datas, hptrs[0], strides, fmts, True, stream)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        st = dec.decode_into(datas, hptrs[0], strides, fmts, True, stream)
    torch.cuda.synchronize(dev)
    cabi_s = time.perf_counter() - t0
    assert all(s == 0 for s in st)
    if rank == 0:
        assert torch.equal(host_sets[0][:out_bytes], dev_sets[0][:out_bytes].cpu()), "host-output path differs from device-output path"
    dec.close()

    # ---- through the unchanged nvimgcodec API: the reference framework dlopen()s libb200jpeg_ext.so from the extensions
    # directory (src/plugin_framework.cpp:281-351); GPU_ONLY backends, so only the plugin can have decoded ----
    sess = ref.Session(ext_dir=nb.EXT_DIR, with_ljt=False)
    fdec = [sess.decoder(device_id=local_rank, backends=ref.BACKEND_GPU_ONLY, options=opts) for _ in range(2)]
    fdims, ffmts = dims, [ref.I_RGB] * n
    for t in dev_sets:
        t.zero_()
    torch.cuda.synchronize(dev)

    def through_api(rounds, depth, d2h, kind=ref.BUFFER_DEVICE, bufs=None):
        a, b = bufs or (ptrs[0], ptrs[1])
        copy = (host_sets[0].data_ptr(), host_sets[1].data_ptr(), dev_sets[0].data_ptr(), dev_sets[1].data_ptr(), out_bytes) if d2h else None
        st, secs = fdec[0].decode_pipelined(datas, a, b, fdims, ffmts, strides, kind, stream, rounds=rounds, depth=depth, d2h=copy,
                                            other=fdec[1] if depth > 1 else None)
        assert all(x == ref.STATUS_SUCCESS for x in st), st
        return secs

    through_api(2, 2, True)
    torch.cuda.synchronize(dev)
    if want is not None:  # what nvimgcodecDecoderDecode delivered, on the device and after the copy to the host
        for t in (dev_sets[0].cpu().numpy(), dev_sets[1].cpu().numpy(), host_sets[0].numpy(), host_sets[1].numpy()):
            for i in (0, n // 2, n - 1):
                w, h = dims[i]
                if not np.array_equal(t[offs[i]:offs[i + 1]].reshape(h, w, 3), want[i]):
                    raise SystemExit("PARITY FAILURE through nvimgcodecDecoderDecode (image %d)" % i)
    api = {}
    steps = max(2, args.steps)
    for key, depth, d2h in (("dev_serial", 1, False), ("dev_pipelined", 2, False), ("host_serial", 1, True), ("host_pipelined", 2, True)):
        through_api(2, depth, d2h)
        barrier()
        api[key] = through_api(steps, depth, d2h) / steps
    # the user hands HOST buffers to the GPU decoder: the framework stages every image in a temporary device buffer it
    # allocates per call and copies back on the image's stream (src/work.h:120-190)
    through_api(1, 1, False, ref.BUFFER_HOST, (hptrs[0], hptrs[1]))
    barrier()
    hb_steps = max(2, min(5, args.steps))
    api["framework_host_buffers"] = through_api(hb_steps, 1, False, ref.BUFFER_HOST, (hptrs[0], hptrs[1])) / hb_steps
    for x in fdec:
        x.close()
    sess.close()

    # ---- secondary reference (north star: "the nvjpeg_ext numbers are added as a secondary reference where that extension
    # installs offline"): the reference's own GPU plugin for this path, built unmodified by `make -C oracle nvjpeg`, through the
    # same framework call with the same bytes and the same device buffers.  N = 1 only; context, not a parity partner (nvJPEG's
    # IDCT is not libjpeg-turbo's). ----
    nvjpeg_ext = None
    nvj_dir = os.path.join(ROOT, "oracle", "_ref", "nvjpeg")
    if rank == 0 and world == 1 and not args.skip_cpu_baseline and not sharded:
        if not os.path.exists(os.path.join(nvj_dir, "libnvjpeg_ext.so")):
            nvjpeg_ext = {"unavailable": "oracle/_ref/nvjpeg/libnvjpeg_ext.so not built (make -C oracle nvjpeg)"}
        else:
            nvjpeg_ext = {"path": "nvimgcodecDecoderDecode -> the reference's nvjpeg_ext (libnvjpeg 12.4 of the CUDA toolkit, static), pageable host "
                                  "JPEG bytes in, device pixels out: the call pattern of e2e_device_out.serial", "backends": {}}
            nsess = ref.Session(ext_dir=nvj_dir, with_ljt=False)
            for name, mask in (("hybrid_cpu_gpu", ref.BACKEND_HYBRID), ("hw_gpu_only", ref.BACKEND_HW)):
                try:
                    nd = nsess.decoder(device_id=local_rank, backends=mask, options=":fancy_upsampling=1")
                    st, _ = nd.decode_pipelined(datas, ptrs[0], ptrs[1], fdims, ffmts, strides, ref.BUFFER_DEVICE, stream, rounds=2, depth=1)
                    torch.cuda.synchronize(dev)
                    if not all(x == ref.STATUS_SUCCESS for x in st):
                        nvjpeg_ext["backends"][name] = {"unavailable": "no decoder of this kind took the images on this device (status 0x%x)" % st[0]}
                        nd.close()
                        continue
                    entry = {}
                    if want is not None:
                        got = dev_sets[0].cpu().numpy()
                        d0 = np.abs(got[offs[0]:offs[1]].astype(np.int16) - want[0].reshape(-1).astype(np.int16))
                        entry["vs_libjpeg_turbo_pixels"] = {"max_abs_diff": int(d0.max()), "identical_fraction": float((d0 == 0).mean())}
                    nsteps = max(3, min(10, args.steps))
                    st, secs = nd.decode_pipelined(datas, ptrs[0], ptrs[1], fdims, ffmts, strides, ref.BUFFER_DEVICE, stream, rounds=nsteps, depth=1)
                    torch.cuda.synchronize(dev)
                    entry.update({"value": pixels / 1e6 / (secs / nsteps), "unit": "Mpixel/s", "ms_per_step": secs / nsteps * 1e3, "steps": nsteps})
                    nvjpeg_ext["backends"][name] = entry
                    nd.close()
                except Exception as e:  # noqa: BLE001  (a secondary reference must not take the bench down)
                    nvjpeg_ext["backends"][name] = {"unavailable": "%s: %s" % (type(e).__name__, e)}
            nsess.close()

    # ---- the floor of any host->host number on this box: the raw device -> pinned host copy of one batch of pixels, every rank
    # at once (at 8 GPUs the box's aggregate device -> host rate, not the decoder, bounds e2e) ----
    barrier()
    floor_steps = max(3, min(10, args.steps))
    for _ in range(2):
        host_sets[0][:out_bytes].copy_(dev_sets[0][:out_bytes], non_blocking=True)
    barrier()
    t0 = time.perf_counter()
    for _ in range(floor_steps):
        host_sets[0][:out_bytes].copy_(dev_sets[0][:out_bytes], non_blocking=True)
    torch.cuda.synchronize(dev)
    api["copy_floor"] = (time.perf_counter() - t0) / floor_steps

    # ---- the floors of the host bytes -> device pixels number, every rank at once: (1) the raw pinned host -> device copy of one
    # batch of entropy-coded bytes, (2) the pageable -> pinned staging copy of the same bytes on the threads this rank owns
    # (what the decoder's StagePool does before its upload): at N > 1 the ranks share the box's PCIe root and memory bus ----
    import ctypes
    from concurrent.futures import ThreadPoolExecutor
    scan_total = int(stats["scan_bytes"])
    pin_in = torch.empty(scan_total + 64, dtype=torch.uint8).pin_memory()
    dev_in = torch.empty(scan_total + 64, dtype=torch.uint8, device=dev)
    for _ in range(2):
        dev_in.copy_(pin_in, non_blocking=True)
    barrier()
    t0 = time.perf_counter()
    for _ in range(floor_steps):
        dev_in.copy_(pin_in, non_blocking=True)
    torch.cuda.synchronize(dev)
    api["h2d_floor"] = (time.perf_counter() - t0) / floor_steps
    pageable = np.frombuffer(b"".join(datas), dtype=np.uint8)[:scan_total].copy()
    nthreads = max(1, min(8, len(os.sched_getaffinity(0))))
    piece = (scan_total + nthreads - 1) // nthreads  # one piece per thread: a python thread pool is too slow for 256 KB pieces
    jobs = [(o, min(piece, scan_total - o)) for o in range(0, scan_total, piece)]
    src0, dst0 = pageable.ctypes.data, pin_in.data_ptr()

    def stage_all(pool):
        list(pool.map(lambda j: ctypes.memmove(dst0 + j[0], src0 + j[0], j[1]), jobs))
    with ThreadPoolExecutor(nthreads) as pool:
        stage_all(pool)
        barrier()
        t0 = time.perf_counter()
        for _ in range(floor_steps):
            stage_all(pool)
        api["stage_floor"] = (time.perf_counter() - t0) / floor_steps
    del pin_in, dev_in

    # ---- reduce over ranks (max time) ----
    keys = ["dev_serial", "dev_pipelined", "host_serial", "host_pipelined", "framework_host_buffers", "copy_floor", "h2d_floor", "stage_floor"]
    t = torch.tensor([ms_total, cabi_s * 1e3 / args.steps] + [api[k] * 1e3 for k in keys] + [float(pixels), float(out_bytes), float(n),
                     float(stats["scan_bytes"])], dtype=torch.float64, device=dev)
    per_rank_ms = [ms_total / args.steps]
    tot = t.clone()
    if world > 1:
        every = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(every, t)
        per_rank_ms = [float(x[0]) / args.steps for x in every]  # the spread behind the max (reported in config)
        tot = torch.stack(every).sum(0)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, cabi_ms = float(t[0]), float(t[1])
    api_ms = {k: float(t[2 + i]) for i, k in enumerate(keys)}
    job_pixels, job_out_bytes, job_images, job_scan = float(tot[2 + len(keys)]), float(tot[3 + len(keys)]), float(tot[4 + len(keys)]), float(tot[5 + len(keys)])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    mpix_step = job_pixels / 1e6
    ms_step = ms_total / args.steps
    value = mpix_step / (ms_step / 1e3)
    peak, peak_kind = hbm_peak()
    # algorithmic bytes per stage for this rank's batch: compressed bytes S (read by every Huffman pass and once more
    # by the reconstruction kernel, which decodes the AC symbols itself), per block 2 B DC + 4 B bit position, planes 64 B/block,
    # output bytes
    S, blocks, out_b, R = stats["scan_bytes"], stats["blocks"], stats["out_bytes"], stats["record_bytes"]
    alg = {"unstuff": 2 * S, "huff_sync": 2 * S, "huff_tail": S * stats["tail_redone"] // max(1, stats["subsequences"]),
           "huff_write": S + 6 * blocks, "dc_scan": 4 * blocks, "idct": S + 6 * blocks + 64 * blocks, "color": 64 * blocks + out_b}
    kernels = {"unstuff": "k_unstuff_count + k_unstuff_scan + k_unstuff_write", "huff_sync": "k_huff_sync x2 (rounds 0 and 1)",
               "huff_tail": "k_huff_tail_round + k_huff_sync_tail (work-list rounds)", "huff_write": "k_subseq_scan + k_huff_write (block pass)",
               "dc_scan": "k_dc_scan", "idct": "k_idct (AC Huffman decode + dequant + IDCT)", "color": "k_color"}
    traffic = {}
    try:  # DRAM bytes per stage from the committed ncu --set full capture of the same 64-image workload
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            tj = json.load(f)
        if tj.get("workload_images") == n and args.workload == "decode":
            traffic = {k: v["dram_bytes"] for k, v in tj["stages"].items()}
    except Exception:
        pass
    stage_ms = {k: v / args.steps for k, v in stage_acc.items()}
    stages = []
    for k in Decoder.STAGES:
        if k in stage_ms and stage_ms[k] > 0:
            ach = alg[k] / (stage_ms[k] * 1e-3) / 1e9
            stages.append({"stage": k, "kernel": kernels[k], "ms": stage_ms[k], "algorithmic_bytes": alg[k], "achieved": ach, "frac": ach / peak,
                           "traffic": traffic.get(k), "traffic_over_algorithmic": (traffic[k] / alg[k]) if traffic.get(k) and alg[k] else None})
    roof = None
    if stages:
        dom = max(stages, key=lambda x: x["ms"])  # the stage with the largest CUDA-event time, no exclusions
        huff_ms = stage_ms.get("huff_sync", 0) + stage_ms.get("huff_tail", 0) + stage_ms.get("huff_write", 0)
        roof = {"bound": "hbm", "kernel": dom["kernel"], "stage": dom["stage"], "achieved": dom["achieved"], "peak": peak, "peak_kind": peak_kind,
                "unit": "GB/s", "frac": dom["frac"], "traffic": dom["traffic"], "algorithmic_bytes": dom["algorithmic_bytes"], "ms": dom["ms"],
                "pipeline": {"algorithmic_bytes": S + out_b, "achieved": (S + out_b) / (ms_step * 1e-3) / 1e9,
                             "frac": (S + out_b) / (ms_step * 1e-3) / 1e9 / peak,
                             "single_stream_ms": sum(x["ms"] for x in stages)},
                "stages": stages,
                "huffman_stage": {"ms": huff_ms, "algorithmic_bytes": S + 6 * blocks,
                                  "frac": (S + 6 * blocks) / (huff_ms * 1e-3) / 1e9 / peak if huff_ms else None,
                                  "what": "k_huff_sync x2 + work-list rounds + k_subseq_scan + k_huff_write; the AC symbols are decoded "
                                          "a last time inside k_idct"},
                "huffman_gbit_per_s": S * 8 / (huff_ms * 1e-3) / 1e9 if huff_ms else None}

    def rate(ms):
        return {"value": mpix_step / (ms / 1e3), "unit": "Mpixel/s", "images_per_s": job_images / (ms / 1e3), "ms_per_step": ms}

    # the copy of batch k overlaps the decode of batch k + 1 in both call patterns (the copies run on their own stream); the
    # headline is the faster one, both are reported
    best = "host_pipelined" if api_ms["host_pipelined"] <= api_ms["host_serial"] else "host_serial"
    e2e = rate(api_ms[best])
    e2e.update({"h2d_bytes_per_step": int(job_scan), "d2h_bytes_per_step": int(job_out_bytes), "pattern": "two decode calls in flight" if best == "host_pipelined" else "one decode call at a time",
                "path": "nvimgcodecDecoderDecode + nvimgcodecFutureWaitForAll of the unmodified framework with libb200jpeg_ext.so found in its "
                        "extensions directory (GPU_ONLY backends), pageable host JPEG bytes in, device batch tensor out, one cudaMemcpyAsync "
                        "to a pinned host batch tensor per batch on a copy stream (it overlaps the next batch's decode); steady state over "
                        "`steps` batches, wall clock, streams synchronised at the end",
                "copy_floor_ms": api_ms["copy_floor"], "frac_of_copy_floor": api_ms["copy_floor"] / api_ms[best],
                "two_calls_in_flight": rate(api_ms["host_pipelined"]),
                "copy_floor": "one cudaMemcpyAsync of the batch's %d output bytes per rank, device -> pinned host, all ranks at once: %.1f GB/s per rank"
                              % (int(job_out_bytes / max(1, world)), job_out_bytes / max(1, world) / api_ms["copy_floor"] / 1e6),
                "serial": rate(api_ms["host_serial"]),
                "framework_host_buffers": rate(api_ms["framework_host_buffers"]),
                "c_abi": rate(cabi_ms)})
    e2e["serial"]["path"] = "one decode call at a time: Decode -> WaitForAll -> enqueue the copy -> next Decode"
    e2e["two_calls_in_flight"]["path"] = "two decoders, ping-pong (the framework promises one batch per decoder): Decode(k + 1) before WaitForAll(k)"
    e2e["framework_host_buffers"]["path"] = ("STRIDED_HOST (pinned) target buffers handed to the GPU decoder: the framework allocates a temporary "
                                             "device buffer per image per call and copies back itself (src/work.h:120-190)")
    e2e["c_abi"]["path"] = "b200jpeg_decode_batch with on_host outputs (no framework): pageable JPEG bytes in, pinned host batch tensor out"
    e2e_dev = rate(api_ms["dev_pipelined"])
    e2e_dev.update({"h2d_bytes_per_step": int(job_scan), "d2h_bytes_per_step": 0, "serial": rate(api_ms["dev_serial"]),
                    "h2d_floor_ms": api_ms["h2d_floor"], "stage_floor_ms": api_ms["stage_floor"],
                    "floors": "per rank, all ranks at once: raw cudaMemcpyAsync of the batch's %d entropy-coded bytes pinned host -> device "
                              "(h2d_floor_ms), and the pageable -> pinned copy of the same bytes, one piece per thread on up to 8 of the threads "
                              "this rank may run on (stage_floor_ms); the decoder does the second, then the first, then %.2f ms of kernels"
                              % (int(job_scan / max(1, world)), ms_step),
                    "path": "as e2e without the copy back: host JPEG bytes -> device pixels, what python/decoder.cpp:148-225 hands its users"})
    line = {
        "metric": "decoded_mpixel_per_s", "value": value, "unit": "Mpixel/s", "images_per_s": job_images / (ms_step / 1e3),
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "strong" if sharded else "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(args),
        "details": {"jpeg_kb_avg": S / n / 1e3, "per_gpu_images": n,
                    "l2": "working set per step %.0f MB (stream+planes+output) > 126 MB L2, no explicit flush" % ((2 * S + 70 * blocks + out_b) / 1e6),
                    "host_affinity": numa, "per_rank_ms_per_step": [round(x, 4) for x in per_rank_ms], "sync_rounds": stats["sync_rounds"],
                    "subsequences": stats["subsequences"], "tail_redone": stats["tail_redone"],
                    "sharding": "batch dealt over the ranks by shard_batch (cost = 8 x bytes + pixels), no collective" if sharded
                                else "every rank decodes its own copy of the batch",
                    "parity": "bit-exact vs reference libjpeg_turbo_ext on all %d images before timing (C ABI), and on samples of what "
                              "nvimgcodecDecoderDecode delivered" % n},
        "clocks": clocks.summary(),
        "e2e": e2e,
        "e2e_device_out": e2e_dev,
        "gpu_launches": int(launches),
        "roofline": roof,
        "cpu_baseline": cpu_baseline,
    }
    if nvjpeg_ext is not None:
        line["nvjpeg_ext"] = nvjpeg_ext
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
