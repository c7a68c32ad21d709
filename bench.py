"""bench.py -- images/s of the predict path on synthetic uint8 images.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--model ResNet50|InceptionV3] [--impl b200|reference]
                  [--scaling weak|strong] [--sustain-seconds S] [--no-second-model] [--micro-batch M]

One "step" = one pass of the hot path (K1 resize/preprocess -> all layers -> dense + softmax + top-5) over one
batch per GPU: ResNet50 batch 256 (BASELINE.json configs[2]) or InceptionV3 batch 128 (configs[3]).
  weak scaling (default): every rank processes its own full batch; value = images of all ranks / max-over-ranks time.
  strong scaling        : ONE batch is image-sharded, rank g takes images [g*B/G, (g+1)*B/G) (configs[3]).
`value` is timed with the inputs resident in HBM; `e2e` goes through the public engine call with pinned HOST buffers
(H2D + kernels + D2H inside the timed region).  `sustained` repeats the device-resident step for >= S seconds with the
clocks sampled every 10 ms (the K-step headline lasts only tens of milliseconds, i.e. at boost clocks).  After the
timed loops (never inside them) one step's top-5 / softmax is compared with the CPU oracle on 16 of the batch's
images: `parity_checked`.  With the default model the same measurements are repeated for InceptionV3 batch 128 in the
same process and attached under "inceptionv3".
--impl reference times the CPU restatement of the reference's Keras loop (oracle/, batch-1 predict per image) on the
host cores; that arm imports nothing of the product.
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
CONV_FLOP = {"ResNet50": 7_711_850_496, "InceptionV3": 11_422_336_192}  # per image, SURVEY.md section 8(d)
RING = 4  # distinct input batches cycled through: 4 x 38.5 MB (ResNet50) > 126 MB L2
PARITY_IMAGES = 16


def synthetic_images(model, n, seed):
    size = 224 if model == "ResNet50" else 299
    rng = np.random.default_rng(seed)
    return [rng.integers(0, 256, size=(size, size, 3), dtype=np.uint8) for _ in range(n)]


class ClockSampler(threading.Thread):
    """Samples SM clock, power and throttle reasons during a timed region (NVML, every 10 ms)."""

    def __init__(self, index, period=0.01):
        super().__init__(daemon=True)
        self.index, self.period, self.stop_flag = index, period, False
        self.sm, self.power, self.reasons, self.max_sm = [], [], set(), None
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
                self.power.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(self.period)

    def finish(self):
        self.stop_flag = True
        self.join()
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None,
                "sm_mhz_min": float(np.min(self.sm)) if self.sm else None, "sm_max_mhz": self.max_sm,
                "power_w_median": float(np.median(self.power)) if self.power else None,
                "power_w_max": float(np.max(self.power)) if self.power else None,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


def cpu_reference_rate(model, seconds_budget, steps=None, warmup=0):
    """The reference's loop shape on the host cores: per image, preprocess + batch-1 forward + top-5.
    Oracle code only (no product import): this is also the --impl reference arm."""
    import torch
    from oracle import models_ref, weights as oweights
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    W = oweights.generate(model)
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
        return n / dt, cores, f"{n} synthetic images, batch-1 loop, {dt:.1f} s", dt / 1, n
    per_step = int(max(1, min(BATCH[model], seconds_budget / ((steps + warmup) * max(t_img, 1e-3)))))
    sel = [imgs[i % 64] for i in range(per_step)]
    for _ in range(warmup):
        models_ref.predict_arrays(model, sel, W, batch=1)
    t0 = time.perf_counter()
    for _ in range(steps):
        models_ref.predict_arrays(model, sel, W, batch=1)
    dt = time.perf_counter() - t0
    return per_step * steps / dt, cores, f"{per_step} of {BATCH[model]} images per step, batch-1 loop", dt / steps, per_step


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    model = args.model
    rate, cores, sample, step_s, per_step = cpu_reference_rate(model, 120.0, steps=args.steps, warmup=args.warmup)
    cfg = workload_config(model, args, BATCH[model])
    cfg["reference_sample"] = (f"each step runs {per_step} of the {BATCH[model]} images of the batch through the reference's "
                               "batch-1 loop (a bounded sample; the rate is per image)")
    line = {"impl": "reference", "metric": "images_per_second", "value": rate, "unit": "images/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_s * 1e3, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
            "cpu_baseline": {"value": rate, "unit": "images/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": rate, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "CPU restatement of the reference's Keras loop (TensorFlow/Keras unavailable offline), torch CPU fp32"}
    print(json.dumps(line))


def workload_config(model, args, n_local):
    size = 224 if model == "ResNet50" else 299
    return {"workload": f"{model} synthetic uint8 {size}x{size}x3, batch {BATCH[model]} "
                        f"{'per GPU' if args.scaling == 'weak' else 'image-sharded over the GPUs'}, "
                        f"resize+preprocess+predict+top5 (BASELINE.json configs[{2 if model == 'ResNet50' else 3}])",
            "batch_per_gpu": n_local, "micro_batch": args.micro_batch,
            "l2": f"ring of {RING} distinct input batches ({RING * n_local * size * size * 3 / 1e6:.0f} MB) cycled; "
                  f"activations stream > 126 MB per step",
            "weights": "fixed-seed random init, Keras layout (ImageNet weights unavailable offline)"}


def measure(model, args, ctx, headline):
    """All measurements of one model on this rank's GPU.  Returns the result dict on rank 0 (None elsewhere)."""
    import torch
    import torch.distributed as dist
    world, rank, local, dev = ctx["world"], ctx["rank"], ctx["local"], ctx["dev"]
    engine_mod = importlib.import_module(PKG + ".engine")
    fe = importlib.import_module(PKG + ".frontend")
    kw = importlib.import_module(PKG + ".keras_weights")
    par = importlib.import_module(PKG + ".parallel")
    B = BATCH[model]
    if args.scaling == "strong":
        a, b = par.shard_range(B, world, rank)
        n, img_lo = b - a, a
        total_images = B
    else:
        n, img_lo = B, 0
        total_images = B * world

    # One weight replica per GPU: rank 0 folds + packs, the packed blob is broadcast (the only collective).
    eng = engine_mod.Engine(model, device=local, max_batch=n, micro_batch=args.micro_batch, load=False)
    W = None
    if rank == 0:
        W = kw.generate(model)
        eng.load_weights(W)
    if world > 1:
        dist.broadcast(eng.packed_weights(), src=0)
        torch.cuda.synchronize()
        if rank != 0:
            eng.weights_ready()

    ring_host, ring_dev, ring_imgs = [], [], []
    for r in range(RING):
        if args.scaling == "strong":
            imgs = synthetic_images(model, B, seed=r)[img_lo:img_lo + n]  # every rank draws the batch, keeps its slice
        else:
            imgs = synthetic_images(model, n, seed=1000 * rank + r)
        if r == 0:
            ring_imgs = imgs[:PARITY_IMAGES]
        hb = eng.pin(fe.HostBatch(imgs, eng.size))
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

    # ---- device-resident throughput: EXACTLY args.steps steps --------------------
    for i in range(args.warmup):
        eng.infer_ptrs(*ring_dev[i % RING], n)
    barrier()
    if args.ncu_step and headline:
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
    value = total_images * args.steps / (ms_total * 1e-3)

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
    clocks = sampler.finish()
    e2e_ms = max_over_ranks(ev0.elapsed_time(ev1))
    hb = ring_host[0]
    h2d = int(hb.pixels.nbytes + hb.img_off.nbytes + hb.hw.nbytes + hb.xtab.nbytes + hb.ytab.nbytes)
    d2h = int(n * 5 * 8)
    e2e = {"value": total_images * args.steps / (e2e_ms * 1e-3), "unit": "images/s", "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": d2h, "api": "Engine.submit / Engine.collect (two batches in flight)"}

    # ---- the same step for seconds: does the number survive the power cap? --------
    sustained = None
    if args.sustain_seconds > 0:
        burst = max(10, int(0.25 / max(ms_step * 1e-3, 1e-5)))  # ~0.25 s of steps per event pair
        barrier()
        s2 = ClockSampler(local)
        s2.start()
        t_end = time.perf_counter() + args.sustain_seconds
        tot_ms, tot_steps, i = 0.0, 0, 0
        tail_ms, tail_steps = 0.0, 0
        while True:
            ev0.record(eng.stream)
            for _ in range(burst):
                eng.infer_ptrs(*ring_dev[i % RING], n)
                i += 1
            ev1.record(eng.stream)
            ev1.synchronize()
            ms = ev0.elapsed_time(ev1)
            tot_ms += ms
            tot_steps += burst
            stop = time.perf_counter() >= t_end
            if world > 1:  # all ranks leave together
                flag = torch.tensor([1.0 if stop else 0.0], device=dev)
                dist.all_reduce(flag, op=dist.ReduceOp.MAX)
                stop = flag.item() > 0
            if stop:
                tail_ms, tail_steps = ms, burst  # the last burst: the clocks have settled by now
                break
        sc = s2.finish()
        tot_ms, tail_ms = max_over_ranks(tot_ms), max_over_ranks(tail_ms)
        sustained = {"seconds": tot_ms * 1e-3, "steps": tot_steps, "value": total_images * tot_steps / (tot_ms * 1e-3),
                     "ms_per_step": tot_ms / tot_steps, "unit": "images/s",
                     "last_burst": {"steps": tail_steps, "value": total_images * tail_steps / (tail_ms * 1e-3)}, "clocks": sc}

    if rank != 0:
        eng.close()
        return None

    # ---- parity of one step of this very plan against the CPU oracle (outside every timed region) --------
    parity = {"checked": False}
    try:
        from oracle import models_ref
        eng.upload(ring_host[0])
        eng.infer_device(n, want_probs=True)
        eng.sync()
        k = min(PARITY_IMAGES, n)
        probs = eng.d_probs[:k].cpu().numpy()
        idx = eng.d_idx[:k].cpu().numpy()
        ref_probs, ref_logits = models_ref.predict_arrays(model, ring_imgs[:k], W, batch=k)
        dsm = float(np.abs(probs - ref_probs).max())
        own_top5 = all(list(np.argsort(probs[r], kind="stable")[-5:][::-1]) == list(idx[r]) for r in range(k))
        parity = {"checked": True, "images": k, "softmax_max_abs": dsm, "tolerance": 2e-2,
                  "top1_equal": int((idx[:, 0] == ref_probs.argmax(1)).sum()), "top5_consistent_with_softmax": bool(own_top5),
                  "passed": bool(dsm <= 2e-2 and own_top5 and np.isfinite(probs).all())}
    except Exception as exc:  # the bench line still prints; the flag says the check did not run
        parity = {"checked": False, "error": f"{type(exc).__name__}: {exc}"}

    # ---- roofline of the dominant kernel (per-launch CUDA events) -----------
    micro = args.micro_batch if args.micro_batch > 0 else eng.max_batch
    waves = (n + micro - 1) // micro
    eng.preprocess(n)  # input tensor of the last uploaded batch
    eng.profile_forward(n)
    ms = np.minimum.reduce([eng.profile_forward(n) for _ in range(3)])
    ops = eng.ops()
    kinds = {0: "conv", 1: "maxpool", 2: "avgpool", 3: "gap", 4: "dense", 5: "softmax_top5"}
    per_kind = {}
    for (name, kind, impl, macs, ab, wb), t in zip(ops, ms):
        if kind == 0 and macs == 0:
            continue  # a conv computed inside another op's launch (merged / chained): reported by that op
        key = kinds[kind] + ("_tcgen05" if kind == 0 and impl == 1 else ("_cuda_core" if kind == 0 else ""))
        d = per_kind.setdefault(key, {"ms": 0.0, "flop": 0.0, "bytes": 0.0, "launch_sites": 0})
        d["ms"] += float(t)
        d["flop"] += 2.0 * macs * n
        d["bytes"] += float(ab) * n + float(wb) * waves
        d["launch_sites"] += 1 if t > 0 else 0
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    tf_peak = peaks.get("bf16_tflops_sustained", 1400.0)
    src = "measured (MEASURED_PEAKS.json, sustained)" if peaks else "fallback (B200_PROFILING.md, sustained ~1.4 PF)"
    # DRAM bytes per launch of the dominant kernel class from the committed ncu capture of this same
    # step (profiles/*_launches_final*_summary.csv; not re-measured here: ncu cannot run inside a timed bench)
    traffic = None
    try:
        import csv as _csv
        import glob as _glob
        pat = "r*_launches_final_summary.csv" if model == "ResNet50" else "r*_launches_inceptionv3_summary.csv"
        f = sorted(_glob.glob(os.path.join(ROOT, "profiles", pat)))[-1]
        rows = [r for r in _csv.DictReader(l for l in open(f) if not l.startswith("#")) if r["kernel"].startswith("conv")]
        if rows:
            traffic = sum(float(r["dram_read_MB"]) + float(r["dram_write_MB"]) for r in rows) * 1e6 / sum(int(r["launches"]) for r in rows)
    except Exception:
        traffic = None
    dom = max(per_kind, key=lambda k: per_kind[k]["ms"])
    d = per_kind[dom]
    step_flop = CONV_FLOP[model] * n
    if dom.startswith("conv"):
        ach = d["flop"] / (d["ms"] * 1e-3) / 1e12
        roof = {"kernel": dom, "bound": "tensor", "achieved": ach, "peak": tf_peak, "unit": "TFLOP/s",
                "frac": ach / tf_peak, "traffic": traffic, "traffic_unit": "DRAM bytes per launch (ncu, profiles/)",
                "launches_per_step": d["launch_sites"] * waves,
                "algorithmic_flop_per_launch": d["flop"] / max(1, d["launch_sites"]), "peak_source": src,
                "share_of_step": d["ms"] / float(ms.sum()),
                # the whole step (all kernels, PDL overlap included) against the same peak
                "frac_of_step": step_flop / (ms_step * 1e-3) / 1e12 / tf_peak if args.scaling == "weak" else None,
                "frac_sustained": (step_flop / (sustained["ms_per_step"] * 1e-3) / 1e12 / tf_peak) if sustained else None}
    else:
        bw_peak = peaks.get("hbm_gbs", 6650.0)
        ach = d["bytes"] / (d["ms"] * 1e-3) / 1e9
        roof = {"kernel": dom, "bound": "hbm", "achieved": ach, "peak": bw_peak, "unit": "GB/s", "frac": ach / bw_peak,
                "traffic": None, "peak_source": src, "share_of_step": d["ms"] / float(ms.sum())}
    for k, d in per_kind.items():
        d["tflops"] = d["flop"] / (d["ms"] * 1e-3) / 1e12 if d["ms"] > 0 else 0.0
        d["gbs"] = d["bytes"] / (d["ms"] * 1e-3) / 1e9 if d["ms"] > 0 else 0.0
        d["hbm_frac"] = d["gbs"] / peaks.get("hbm_gbs", 6650.0)
    if args.ops_csv and headline:
        with open(args.ops_csv, "w") as fh:
            fh.write("op,kind,impl,ms,gflop,mbytes,tflops,gbs\n")
            for (name, kind, impl, macs, ab, wb), t in zip(ops, ms):
                fl, by = 2.0 * macs * n, float(ab) * n + wb
                fh.write(f"{name},{kinds[kind]},{impl},{t:.5f},{fl / 1e9:.3f},{by / 1e6:.3f},"
                         f"{fl / max(t, 1e-9) / 1e9:.2f},{by / max(t, 1e-9) / 1e6:.1f}\n")
    eng.close()
    return {"value": value, "ms_per_step": ms_step, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
            "roofline": roof, "kernels": per_kind, "sustained": sustained, "parity": parity,
            "parity_checked": bool(parity.get("checked") and parity.get("passed")),
            "config": workload_config(model, args, n)}


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
    ctx = {"world": world, "rank": rank, "local": local, "dev": torch.device("cuda", local)}
    model = args.model
    head = measure(model, args, ctx, headline=True)
    second = None
    if model == "ResNet50" and not args.no_second_model:
        second = measure("InceptionV3", args, ctx, headline=False)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        rate, cores, sample, _, _ = cpu_reference_rate(model, 15.0)
        cpu = {"value": rate, "unit": "images/s", "cores": cores, "kind": "port", "sample": sample}

    if rank == 0:
        line = {"metric": "images_per_second", "value": head["value"], "unit": "images/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": args.scaling,
                "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": head["config"],
                "e2e": head["e2e"], "gpu_launches": head["gpu_launches"], "clocks": head["clocks"], "roofline": head["roofline"],
                "cpu_baseline": cpu, "parity_checked": head["parity_checked"], "parity": head["parity"],
                "sustained": head["sustained"], "kernels": head["kernels"]}
        if second is not None:
            line["inceptionv3"] = {k: second[k] for k in ("value", "ms_per_step", "e2e", "gpu_launches", "clocks", "roofline",
                                                          "parity_checked", "parity", "sustained", "kernels", "config")}
            line["inceptionv3"]["unit"] = "images/s"
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--model", default="ResNet50", choices=list(BATCH))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--sustain-seconds", type=float, default=2.0)
    ap.add_argument("--no-second-model", action="store_true", help="ResNet50 only (skip the attached InceptionV3 measurement)")
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
