"""bench.py -- images/s of the predict path on synthetic uint8 images.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--model ResNet50|InceptionV3]
                  [--impl b200|reference] [--micro-batch M]

One "step" = one pass of the hot path (K1 resize/preprocess -> all layers ->
dense + softmax + top-5) over one batch per GPU: ResNet50 batch 256 (BASELINE.json
configs[2]) or InceptionV3 batch 128 (configs[3]).  Weak scaling: every rank
processes its own batch; value = images all ranks processed / max-over-ranks time.
`value` is timed with the inputs resident in HBM; `e2e` goes through the public
engine call with pinned HOST buffers (H2D + kernels + D2H inside the timed region).
--impl reference times the CPU restatement of the reference's Keras loop
(oracle/, batch-1 predict per image) on the host cores.
"""
import argparse
import importlib
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "distributed-machine-learning_b200"
BATCH = {"ResNet50": 256, "InceptionV3": 128}
RING = 4  # distinct input batches cycled through: 4 x 38.5 MB (ResNet50) > 126 MB L2


def synthetic_images(model, n, seed):
    size = 224 if model == "ResNet50" else 299
    rng = np.random.default_rng(seed)
    return [rng.integers(0, 256, size=(size, size, 3), dtype=np.uint8) for _ in range(n)]


class ClockSampler(threading.Thread):
    """Samples SM clocks and throttle reasons during the timed region (NVML)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.sm, self.reasons, self.max_sm = index, False, [], set(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_sm = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
                 nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake"}
        while not self.stop_flag:
            try:
                self.sm.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.05)

    def result(self):
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_sm,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


def cpu_reference_rate(model, seconds_budget, steps=None, warmup=0):
    """The reference's loop shape on the host cores: per image, preprocess + batch-1 forward + top-5."""
    import torch
    from oracle import models_ref
    kw = importlib.import_module(PKG + ".keras_weights")
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    W = kw.generate(model)
    imgs = synthetic_images(model, 64, 0)
    models_ref.predict_arrays(model, imgs[:2], W, batch=1)  # warm-up (thread pool, oneDNN primitives)
    t0 = time.perf_counter()
    models_ref.predict_arrays(model, imgs[:6], W, batch=1)
    t_img = (time.perf_counter() - t0) / 6
    if steps is None:
        n = int(max(4, min(2048, seconds_budget / max(t_img, 1e-3))))
        sel = [imgs[i % 64] for i in range(n)]
        t0 = time.perf_counter()
        models_ref.predict_arrays(model, sel, W, batch=1)
        dt = time.perf_counter() - t0
        return n / dt, cores, f"{n} synthetic images, batch-1 loop, {dt:.1f} s", dt / 1
    per_step = int(max(1, min(BATCH[model], seconds_budget / ((steps + warmup) * max(t_img, 1e-3)))))
    sel = [imgs[i % 64] for i in range(per_step)]
    for _ in range(warmup):
        models_ref.predict_arrays(model, sel, W, batch=1)
    t0 = time.perf_counter()
    for _ in range(steps):
        models_ref.predict_arrays(model, sel, W, batch=1)
    dt = time.perf_counter() - t0
    return per_step * steps / dt, cores, f"{per_step} of {BATCH[model]} images per step, batch-1 loop", dt / steps


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    model = args.model
    rate, cores, sample, step_s = cpu_reference_rate(model, 120.0, steps=args.steps, warmup=args.warmup)
    line = {"impl": "reference", "metric": "images_per_second", "value": rate, "unit": "images/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_s * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(model, args),
            "cpu_baseline": {"value": rate, "unit": "images/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": rate, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "CPU restatement of the reference's Keras loop (TensorFlow/Keras unavailable offline), torch CPU fp32"}
    print(json.dumps(line))


def workload_config(model, args):
    size = 224 if model == "ResNet50" else 299
    return {"workload": f"{model} synthetic uint8 {size}x{size}x3, batch {BATCH[model]} per GPU, "
                        f"resize+preprocess+predict+top5 (BASELINE.json configs[{2 if model == 'ResNet50' else 3}])",
            "batch_per_gpu": BATCH[model], "micro_batch": args.micro_batch,
            "l2": f"ring of {RING} distinct input batches ({RING * BATCH[model] * size * size * 3 / 1e6:.0f} MB) cycled; "
                  f"activations stream > 126 MB per step",
            "weights": "fixed-seed random init, Keras layout (ImageNet weights unavailable offline)"}


def run_b200(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    model = args.model
    n = BATCH[model]
    engine_mod = importlib.import_module(PKG + ".engine")
    fe = importlib.import_module(PKG + ".frontend")
    kw = importlib.import_module(PKG + ".keras_weights")

    # One weight replica per GPU: rank 0 folds + packs, the packed blob is broadcast (the only collective).
    eng = engine_mod.Engine(model, device=local, max_batch=n, micro_batch=args.micro_batch, load=False)
    if rank == 0:
        eng.load_weights(kw.generate(model))
    if world > 1:
        dist.broadcast(eng.packed_weights(), src=0)
        torch.cuda.synchronize()
        if rank != 0:
            eng.weights_ready()

    dev = torch.device("cuda", local)
    ring_host, ring_dev = [], []
    for r in range(RING):
        hb = eng.pin(fe.HostBatch(synthetic_images(model, n, seed=1000 * rank + r), eng.size))
        ring_host.append(hb)
        ring_dev.append(tuple(torch.from_numpy(getattr(hb, k)).to(dev) for k in ("pixels", "img_off", "hw", "xtab", "ytab")))
    torch.cuda.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident throughput --------------------------------------
    for i in range(args.warmup):
        eng.infer_ptrs(*ring_dev[i % RING], n)
    barrier()
    if args.ncu_step:
        # exactly one step between cudaProfilerStart/Stop: run under `ncu --profile-from-start off`
        torch.cuda.profiler.start()
        eng.infer_ptrs(*ring_dev[0], n)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    sampler = ClockSampler(local)
    sampler.start()
    l0 = eng.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(eng.stream)
    for i in range(args.steps):
        eng.infer_ptrs(*ring_dev[i % RING], n)
    ev1.record(eng.stream)
    barrier()
    launches = eng.launch_count() - l0
    ms_total = max_over_ranks(ev0.elapsed_time(ev1))
    ms_step = ms_total / args.steps
    value = world * n * args.steps / (ms_total * 1e-3)

    # ---- end to end through the public engine call, host buffers ----------
    # submit()/collect(): two batches in flight, H2D of step i+1 overlaps the kernels of step i; every
    # step's inputs come from pinned host memory and every step's top-5 is read back to the host.
    def e2e_steps(k):
        pending = None
        for i in range(k):
            t = eng.submit(ring_host[i % RING])
            if pending is not None:
                eng.collect(pending)
            pending = t
        if pending is not None:
            eng.collect(pending)

    e2e_steps(min(3, args.warmup))
    barrier()
    ev0.record(eng.stream)
    e2e_steps(args.steps)
    ev1.record(eng.stream)
    barrier()
    sampler.stop_flag = True
    sampler.join()
    e2e_ms = max_over_ranks(ev0.elapsed_time(ev1))
    hb = ring_host[0]
    h2d = int(hb.pixels.nbytes + hb.img_off.nbytes + hb.hw.nbytes + hb.xtab.nbytes + hb.ytab.nbytes)
    d2h = int(n * 5 * 8)
    e2e = {"value": world * n * args.steps / (e2e_ms * 1e-3), "unit": "images/s", "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": d2h}

    # ---- roofline of the dominant kernel (per-launch CUDA events) -----------
    roof = None
    per_kind = {}
    if rank == 0:
        eng.preprocess(n)  # input tensor of the last uploaded batch
        eng.profile_forward(n)
        ms = np.minimum.reduce([eng.profile_forward(n) for _ in range(3)])
        ops = eng.ops()
        kinds = {0: "conv", 1: "maxpool", 2: "avgpool", 3: "gap", 4: "dense", 5: "softmax_top5"}
        for (name, kind, impl, macs, ab, wb), t in zip(ops, ms):
            if kind == 0 and macs == 0:
                continue  # follower of a merged pair of 1x1 convolutions: no launch of its own, reported by its leader
            key = kinds[kind] + ("_tcgen05" if kind == 0 and impl == 1 else ("_cuda_core" if kind == 0 else ""))
            d = per_kind.setdefault(key, {"ms": 0.0, "flop": 0.0, "bytes": 0.0, "launch_sites": 0})
            d["ms"] += float(t)
            d["flop"] += 2.0 * macs * n
            d["bytes"] += float(ab) * n + float(wb) * ((n + eng_micro(eng, args) - 1) // eng_micro(eng, args))
            d["launch_sites"] += 1
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        tf_peak = peaks.get("bf16_tflops_sustained", 1400.0)
        src = "measured (MEASURED_PEAKS.json, sustained)" if peaks else "fallback (B200_PROFILING.md, sustained ~1.4 PF)"
        # DRAM bytes per launch of the dominant kernel class from the committed ncu capture of this same
        # step (profiles/*_launches_final_summary.csv; not re-measured here: ncu cannot run inside a timed bench)
        traffic = None
        try:
            import csv as _csv
            import glob as _glob
            f = sorted(_glob.glob(os.path.join(ROOT, "profiles", "r*_launches_final_summary.csv")))[-1]
            rows = [r for r in _csv.DictReader(l for l in open(f) if not l.startswith("#")) if r["kernel"].startswith("conv_tc")]
            if model == "ResNet50" and rows:
                traffic = sum(float(r["dram_read_MB"]) + float(r["dram_write_MB"]) for r in rows) * 1e6 / sum(int(r["launches"]) for r in rows)
        except Exception:
            traffic = None
        dom = max(per_kind, key=lambda k: per_kind[k]["ms"])
        d = per_kind[dom]
        if dom.startswith("conv"):
            ach = d["flop"] / (d["ms"] * 1e-3) / 1e12
            roof = {"kernel": dom, "bound": "tensor", "achieved": ach, "peak": tf_peak, "unit": "TFLOP/s",
                    "frac": ach / tf_peak, "traffic": traffic, "traffic_unit": "DRAM bytes per launch (ncu, profiles/)",
                    "launches_per_step": d["launch_sites"] * ((n + eng_micro(eng, args) - 1) // eng_micro(eng, args)),
                    "algorithmic_flop_per_launch": d["flop"] / max(1, d["launch_sites"]), "peak_source": src,
                    "share_of_step": d["ms"] / float(ms.sum())}
        else:
            bw_peak = peaks.get("hbm_gbs", 6650.0)
            ach = d["bytes"] / (d["ms"] * 1e-3) / 1e9
            roof = {"kernel": dom, "bound": "hbm", "achieved": ach, "peak": bw_peak, "unit": "GB/s", "frac": ach / bw_peak,
                    "traffic": None, "peak_source": src, "share_of_step": d["ms"] / float(ms.sum())}
        for k, d in per_kind.items():
            d["tflops"] = d["flop"] / (d["ms"] * 1e-3) / 1e12 if d["ms"] > 0 else 0.0
            d["gbs"] = d["bytes"] / (d["ms"] * 1e-3) / 1e9 if d["ms"] > 0 else 0.0
        if args.ops_csv:
            with open(args.ops_csv, "w") as fh:
                fh.write("op,kind,impl,ms,gflop,mbytes,tflops,gbs\n")
                for (name, kind, impl, macs, ab, wb), t in zip(ops, ms):
                    fl, by = 2.0 * macs * n, float(ab) * n + wb
                    fh.write(f"{name},{kinds[kind]},{impl},{t:.5f},{fl / 1e9:.3f},{by / 1e6:.3f},"
                             f"{fl / max(t, 1e-9) / 1e9:.2f},{by / max(t, 1e-9) / 1e6:.1f}\n")

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        rate, cores, sample, _ = cpu_reference_rate(model, 15.0)
        cpu = {"value": rate, "unit": "images/s", "cores": cores, "kind": "port", "sample": sample}

    if rank == 0:
        conv_flop = {"ResNet50": 7_711_850_496, "InceptionV3": 11_422_336_192}[model]
        line = {"metric": "images_per_second", "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": workload_config(model, args),
                "e2e": e2e, "gpu_launches": int(launches), "clocks": sampler.result(), "roofline": roof,
                "cpu_baseline": cpu, "kernels": per_kind,
                "conv_tensor_frac_of_step": value / world * conv_flop / 1e12 / (roof["peak"] if roof and roof["bound"] == "tensor" else 1400.0)}
        print(json.dumps(line))
    eng.close()
    if world > 1:
        dist.destroy_process_group()


def eng_micro(eng, args):
    return args.micro_batch if args.micro_batch > 0 else eng.max_batch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--model", default="ResNet50", choices=list(BATCH))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--micro-batch", type=int, default=0)
    ap.add_argument("--ops-csv", default=None)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ncu-step", action="store_true", help="bracket one extra step with cudaProfilerStart/Stop")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
