#!/usr/bin/env python3
"""bench.py — headline benchmark of the B200 batched CTC decoder (driver contract).

Workload (BASELINE.json configs[1]): Viterbi batch, 1M reads x T=4000 x C=5 float32 posteriors,
qstring on, qscale/qbias, collapse_repeats — per GPU ("weak" scaling: reads are independent, every
rank decodes its own 1M-read shard, no collective on the data path).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME|all]

The headline keys of the one JSON line are the Viterbi workload (BASELINE.json configs[1]); with the default
--workload all the line also carries "workloads": {viterbi_f16, beam5, beam32, duplex} — the other BASELINE
configs, measured one after the other in the same process (each after the previous one's arrays are freed),
every entry with its own value / ms_per_step / roofline / e2e / cpu_baseline / clocks.
One "step" = one pass of the hot path over the whole resident batch (one kernel launch).
  value     rows/s, inputs resident in HBM, device-timed with CUDA events on the launching stream,
            barrier + synchronize on both sides, max over ranks.
  e2e       the same metric through the C ABI host entry point (fctc_viterbi_batch) with pinned HOST
            buffers: host->device and device->host copies inside the timed region.
  roofline  algorithmic bytes per launch / mean launch time, vs the measured HBM copy peak.
  cpu_baseline  the CPU oracle (C++ restatement of the reference; the Rust reference itself cannot be
            built in this image) on the box's host cores, on a bounded sample of the same bytes.
--impl reference prints the reference arm's line (CPU oracle, all host threads, bounded sample).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "fast-ctc-decode_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

METRIC = "posterior rows decoded/s (Viterbi; beam 5 & 32) at 1/2/4/8 B200 vs host CPU"
UNIT = "rows/s"
SEED = 0xC7C00001

WORKLOADS = {
    # name: reads per GPU, rows, labels, decoder, parameters
    "viterbi": dict(reads=1_000_000, T=4000, C=5, kind="viterbi", qstring=1, qscale=0.7, qbias=2.0, collapse=1,
                    cpu_sample=32768, e2e_reads=32768,
                    label="viterbi 1M reads x T=4000 x C=5 f32, qstring qscale=0.7 qbias=2.0, collapse_repeats"),
    # the same decode over binary16 posteriors (10 B/row); not the headline metric
    "viterbi_f16": dict(reads=1_000_000, T=4000, C=5, kind="viterbi", qstring=1, qscale=0.7, qbias=2.0, collapse=1,
                        cpu_sample=32768, e2e_reads=32768, f16=1,
                        label="viterbi 1M reads x T=4000 x C=5 binary16 posteriors (widened exactly to f32), qstring "
                              "qscale=0.7 qbias=2.0, collapse_repeats"),
    "beam5": dict(reads=262_144, T=4000, C=5, kind="beam", beam=5, thr=0.1, collapse=1, cpu_sample=2048,
                  e2e_reads=65536, label="beam_search 262,144 reads x T=4000 x C=5, beam_size=5, beam_cut_threshold=0.1"),
    "beam32": dict(reads=65_536, T=10_000, C=5, kind="beam", beam=32, thr=0.0, collapse=1, cpu_sample=128,
                   e2e_reads=32768, label="beam_search 65,536 reads x T=10,000 x C=5, beam_size=32, beam_cut_threshold=0.0"),
    "duplex": dict(reads=16_384, T=4000, T2=4000, C=5, kind="duplex", beam=5, thr=0.0, collapse=1, band=32, exp_mode=0,
                   cpu_sample=64, e2e_reads=16384,  # the whole configuration: two chunks, the second copied in under the first's kernel
                   label="duplex 2D envelope beam search: 16,384 pairs x T1=T2=4000 x C=5, beam_size=5, band +-32, exact exp "
                         "(exp_mode 0: glibc's expf/log1pf evaluated on the device, results bit-identical to the CPU reference)"),
    # the same search with CUDA's own expf/log1pf (exp_mode 2, the opt-in fast mode; round 1's duplex figure was measured this way)
    "duplex_cuda_exp": dict(reads=16_384, T=4000, T2=4000, C=5, kind="duplex", beam=5, thr=0.0, collapse=1, band=32, exp_mode=2,
                            cpu_sample=64, e2e_reads=16384,
                            label="duplex 2D envelope beam search: 16,384 pairs x T1=T2=4000 x C=5, beam_size=5, band +-32, "
                                  "exp_mode 2 (CUDA expf/log1pf: last-bit differences from a glibc host)"),
}


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    return {"hbm_gbs": 6650.0}, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.gpu), "-lms", "50"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, power, reasons = [], [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(power)),
                "samples": len(sm), "reasons": sorted(reasons)}


def host_threads():
    import oracle
    return max(1, oracle.hardware_threads())


def cpu_pass(w, post_np, n_threads):
    """One pass of the CPU oracle over post_np [n, T, C] (results discarded): (rows, seconds)."""
    import oracle
    t0 = time.perf_counter()
    if w["kind"] == "viterbi":
        oracle.viterbi_batch(post_np, bool(w["qstring"]), w["qscale"], w["qbias"], bool(w["collapse"]),
                             n_threads=n_threads, keep=False)
    elif w["kind"] == "duplex":
        p1, p2, env = post_np
        oracle.beam_search_duplex_batch(p1, p2, env.astype(np.uint64), w["beam"], w["thr"], bool(w["collapse"]),
                                        min(w["exp_mode"], 1) if w["exp_mode"] != 2 else 0, n_threads=n_threads)
        return p1.shape[0] * p1.shape[1], time.perf_counter() - t0
    else:
        oracle.beam_search_batch(post_np, w["beam"], w["thr"], bool(w["collapse"]), n_threads=n_threads, keep=False)
    return post_np.shape[0] * post_np.shape[1], time.perf_counter() - t0


def cpu_rate(w, post_np, n_threads, min_seconds=0.0, max_passes=64):
    """rows/s of the CPU oracle: whole passes over the sample until min_seconds of wall clock are spent."""
    rows = 0
    secs = 0.0
    passes = 0
    while passes == 0 or (secs < min_seconds and passes < max_passes):
        r, dt = cpu_pass(w, post_np, n_threads)
        rows += r
        secs += dt
        passes += 1
    return rows / secs, secs, passes


def band_envelope_np(n, T1, T2, half):
    c = (np.arange(T1) * T2) // max(T1, 1)
    env = np.stack([np.maximum(0, c - half), np.minimum(T2, c + half + 1)], axis=1).astype(np.int64)
    return np.ascontiguousarray(np.broadcast_to(env, (n, T1, 2)))


def synth_host_sample(w, n_reads, first_read=0):
    """The first n_reads reads of rank 0's shard, as host float32 (generated by the device generator
    when a GPU is present so that both arms see identical bytes; numpy restatement otherwise)."""
    if w["kind"] == "duplex":
        import torch
        import fctc_b200
        dec = fctc_b200.Decoder(0)
        st = torch.cuda.current_stream().cuda_stream
        a = torch.empty((n_reads, w["T"], w["C"]), dtype=torch.float32, device="cuda:0")
        b = torch.empty((n_reads, w["T2"], w["C"]), dtype=torch.float32, device="cuda:0")
        dec.synth_posteriors_device(a.data_ptr(), n_reads, w["T"], w["C"], 1, SEED, first_read, st)
        dec.synth_posteriors_device(b.data_ptr(), n_reads, w["T2"], w["C"], 2, SEED, first_read, st)
        torch.cuda.synchronize()
        out = (a.cpu().numpy(), b.cpu().numpy(), band_envelope_np(n_reads, w["T"], w["T2"], w["band"]))
        dec.close()
        return out, "device generator (same bytes as the GPU arm)"
    try:
        import torch
        if torch.cuda.is_available():
            import fctc_b200
            dec = fctc_b200.Decoder(0)
            t = torch.empty((n_reads, w["T"], w["C"]), dtype=torch.float32, device="cuda:0")
            dec.synth_posteriors_device(t.data_ptr(), n_reads, w["T"], w["C"], 0, SEED, first_read,
                                        torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize()
            if w.get("f16"):  # the binary16 workload decodes the widened values of the rounded array
                t = t.to(torch.float16).to(torch.float32)
            out = t.cpu().numpy()
            del t
            dec.close()
            return out, "device generator (same bytes as the GPU arm)"
    except Exception:
        pass
    rng = np.random.default_rng(SEED)
    z = (rng.random((n_reads, w["T"], w["C"]), dtype=np.float32) * 2 - 1) * 3
    e = np.exp(z - z.max(axis=-1, keepdims=True))
    return (e / e.sum(axis=-1, keepdims=True)).astype(np.float32), "numpy softmax(3(2u-1)) (no GPU present)"


def reference_workload(args, w, steps):
    """One workload of the reference arm: the CPU oracle on all host threads, `steps` passes over a bounded sample."""
    n = w["cpu_sample"]
    post, how = synth_host_sample(w, n)
    threads = host_threads()
    small = tuple(x[: max(1, n // 8)] for x in post) if isinstance(post, tuple) else post[: max(1, n // 8)]
    for _ in range(min(max(args.warmup, 0), 3)):
        cpu_pass(w, small, threads)
    t_total = 0.0
    for _ in range(steps):
        _, dt = cpu_pass(w, post, threads)
        t_total += dt
    value = n * w["T"] * steps / t_total
    return {
        "value": value, "unit": UNIT, "steps": steps, "ms_per_step": 1e3 * t_total / steps, "dtype": "f32",
        "config": {"workload": w["label"], "sample": f"{n} reads per step ({how})"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"{n} reads x {w['T']} rows per step, {steps} steps, one read per worker thread; "
                                   "C++ restatement of the reference (Rust toolchain absent)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }


def run_reference(args, names):
    """Reference arm: the reference's CPU algorithm (oracle port; the Rust build is unavailable here)
    on all host threads; every step decodes one bounded sample of the workload.  Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import oracle
    oracle.build()
    head = reference_workload(args, WORKLOADS[names[0]], args.steps)
    line = {"impl": "reference", "metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": head["config"],
            "cpu_baseline": head["cpu_baseline"], "e2e": head["e2e"]}
    if len(names) > 1:
        line["workloads"] = {nm: reference_workload(args, WORKLOADS[nm], max(1, min(args.steps, 3))) for nm in names[1:]}
    print(json.dumps(line))


def bind_to_gpu_numa_node(torch, local_rank, local_world):
    """Pin this rank to the cores of its GPU's NUMA node (shared evenly with the other ranks of that node), so that
    the pinned staging buffers allocated afterwards are first touched next to the GPU's PCIe root."""
    try:
        prop = torch.cuda.get_device_properties(local_rank)
        bus = f"{prop.pci_domain_id:04x}:{prop.pci_bus_id:02x}:{prop.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{bus}/numa_node") as f:
            node = int(f.read().strip())
        allowed = sorted(os.sched_getaffinity(0))
        cores = allowed
        if node >= 0:
            with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
                ids = set()
                for part in f.read().strip().split(","):
                    a, _, b = part.partition("-")
                    ids.update(range(int(a), int(b or a) + 1))
            cores = [c for c in allowed if c in ids] or allowed
        # ranks that land on the same node split its cores (the topology is the same for every rank of the box)
        share = max(1, len(cores) * 1 // max(local_world, 1))
        mine = cores[(local_rank * share) % len(cores):][:share] or cores
        if local_world > 1 and len(mine) >= 2:
            os.sched_setaffinity(0, mine)
        elif node >= 0:
            os.sched_setaffinity(0, cores)
        return {"numa_node": node, "cores": len(os.sched_getaffinity(0))}
    except Exception as e:  # no sysfs entry, no permission: run unbound
        return {"numa_node": None, "cores": len(os.sched_getaffinity(0)), "note": f"unbound ({type(e).__name__})"}


def profile_figures(name):
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(tpath):
        return None
    with open(tpath) as f:
        return json.load(f).get(name)


def measure_workload(name, w, args, steps, ctx):
    """Device-resident value, roofline, end-to-end leg and (rank 0) CPU baseline of one workload."""
    torch, dist, dec = ctx["torch"], ctx["dist"], ctx["dec"]
    rank, local_rank, world, dev = ctx["rank"], ctx["local_rank"], ctx["world"], ctx["dev"]
    N, T, Cn = args.reads or w["reads"], w["T"], w["C"]
    stream = torch.cuda.current_stream().cuda_stream
    stride = (T + 31) // 32 * 32
    warm = max(args.warmup, 3)

    # ---- resident synthetic input: rank r decodes reads [r*N, (r+1)*N) of the global job ----------
    from fctc_b200.sharding import weak_shard_first_read
    f16 = bool(w.get("f16"))
    eb = 2 if f16 else 4  # bytes per posterior
    if f16:  # generate in float32 chunk by chunk, keep the binary16 rounding of every value
        post = torch.empty((N, T, Cn), dtype=torch.float16, device=dev)
        chunk = 65536
        tmp = torch.empty((min(chunk, N), T, Cn), dtype=torch.float32, device=dev)
        for lo in range(0, N, chunk):
            cnt = min(chunk, N - lo)
            dec.synth_posteriors_device(tmp.data_ptr(), cnt, T, Cn, 0, SEED, weak_shard_first_read(N, rank) + lo, stream)
            post[lo:lo + cnt].copy_(tmp[:cnt])
        del tmp
    else:
        post = torch.empty((N, T, Cn), dtype=torch.float32, device=dev)
        dec.synth_posteriors_device(post.data_ptr(), N, T, Cn, 1 if w["kind"] == "duplex" else 0, SEED,
                                    weak_shard_first_read(N, rank), stream)
    cap = N * stride
    labels = torch.empty(cap, dtype=torch.uint8, device=dev)
    path = torch.empty(cap, dtype=torch.int32, device=dev)
    offsets = torch.empty(N, dtype=torch.int64, device=dev)
    lens = torch.empty(N, dtype=torch.int32, device=dev)
    status = torch.empty(N, dtype=torch.int32, device=dev)
    used = torch.zeros(1, dtype=torch.int64, device=dev)
    quals = post2 = env = scratch = None
    if w["kind"] == "viterbi":
        quals = torch.empty(cap, dtype=torch.uint8, device=dev)

        def step():
            dec.viterbi_batch_device(post.data_ptr(), N, T, Cn, w["qstring"], w["qscale"], w["qbias"], w["collapse"],
                                     labels.data_ptr(), quals.data_ptr(), 1, path.data_ptr(), cap, offsets.data_ptr(),
                                     lens.data_ptr(), status.data_ptr(), stream, f16=f16)
    elif w["kind"] == "duplex":
        T2 = w["T2"]
        post2 = torch.empty((N, T2, Cn), dtype=torch.float32, device=dev)
        dec.synth_posteriors_device(post2.data_ptr(), N, T2, Cn, 2, SEED, weak_shard_first_read(N, rank), stream)
        env = torch.from_numpy(band_envelope_np(N, T, T2, w["band"])).to(dev)
        max_window = 2 * w["band"] + 1
        free_b, _ = torch.cuda.mem_get_info(dev)
        nbytes = min(dec.duplex_scratch_bytes(N, T, T2, Cn, w["beam"], max_window, 0), int(free_b * 0.8) // 256 * 256)
        scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)

        def step():
            used.zero_()
            dec.beam_search_duplex_batch_device(post.data_ptr(), post2.data_ptr(), N, T, T2, Cn, env.data_ptr(), max_window,
                                                w["beam"], w["thr"], w["collapse"], w["exp_mode"], labels.data_ptr(), cap,
                                                offsets.data_ptr(), lens.data_ptr(), status.data_ptr(), used.data_ptr(),
                                                scratch.data_ptr(), nbytes, 0, stream)
    else:
        nbytes = dec.beam_scratch_bytes(N, T, Cn, w["beam"], 0)
        scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)

        def step():
            used.zero_()
            dec.beam_search_batch_device(post.data_ptr(), N, T, Cn, w["beam"], w["thr"], w["collapse"], labels.data_ptr(),
                                         path.data_ptr(), cap, offsets.data_ptr(), lens.data_ptr(), status.data_ptr(),
                                         used.data_ptr(), scratch.data_ptr(), nbytes, 0, stream)
    launches_per_step = 1
    torch.cuda.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warm):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.1)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    ev[0].record()
    for i in range(steps):
        step()
        ev[i + 1].record()
    barrier()
    clocks = sampler.stop()
    total_ms = ev[0].elapsed_time(ev[-1])
    step_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]
    if world > 1:
        t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    ms_per_step = total_ms / steps
    rows_all = N * T * world
    value = rows_all / (ms_per_step * 1e-3)

    # ---- algorithmic bytes of one launch: 4*C per row in, per emitted label 1 (label)
    #      + 1 (quality) + 4 (path) out, + 16 B per read (offset, len, status) -------------------------
    emitted = int(lens.sum(dtype=torch.int64).item())
    ok_reads = int((status == 0).sum().item())
    out_per_label = (1 + 4 + (1 if w.get("qstring") else 0))
    alg_bytes = N * T * Cn * eb + emitted * out_per_label + N * 16
    if w["kind"] == "duplex":  # two posterior arrays + the envelope in, labels out
        alg_bytes = N * (T + w["T2"]) * Cn * 4 + N * T * 16 + emitted + N * 16
    mean_launch_ms = float(np.mean(step_ms))
    peaks, peak_src = measured_peaks()
    achieved = alg_bytes / (mean_launch_ms * 1e-3) / 1e9
    traffic = None
    tj = profile_figures(name)
    if tj and "dram_bytes_per_read" in tj:
        traffic = tj["dram_bytes_per_read"] * N
    elif tj and "dram_bytes_per_row" in tj:
        traffic = tj["dram_bytes_per_row"] * N * T
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                "frac": achieved / peaks["hbm_gbs"], "traffic": traffic,
                "traffic_source": (tj or {}).get("source", None) and
                ("NOT measured in this run: per-row DRAM bytes of the committed ncu --set full capture, scaled to this "
                 "launch's rows — " + tj["source"]),
                "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes, "mean_launch_ms": mean_launch_ms,
                "input_stream_gbs": N * T * Cn * eb / (mean_launch_ms * 1e-3) / 1e9}
    if w["kind"] != "viterbi":
        roofline["note"] = ("the searches are bound by instruction issue and dependent-load latency, not by HBM: frac says "
                            "how little of the HBM roof they use; rows/s and the *_ncu_profile figures are the measure")
    if tj and "issue_slots_busy" in tj:
        # from the committed ncu capture under profiles/ (not measured live): how far the kernel is from the issue limit
        roofline["sm_issue_slots_busy_ncu_profile"] = tj["issue_slots_busy"]
        roofline["warp_instructions_per_row_ncu_profile"] = tj["warp_instructions_per_row"]

    res = {"value": value, "unit": UNIT, "steps": steps, "warmup": warm, "ms_per_step": ms_per_step,
           "dtype": "f32 (binary16 input widened exactly)" if f16 else "f32",
           "config": {"workload": w["label"], "reads_per_gpu": N, "rows_per_read": T, "labels": Cn,
                      "input": "device generator, " + ("peaked rows (dominant label, blank w.p. 0.65), two noisy renderings "
                                                       "of one label track" if w["kind"] == "duplex" else
                                                       "softmax(3(2u-1)) rows") + ", seed 0xC7C00001",
                      "l2": f"input {N * T * Cn * eb / 1e9:.1f} GB per GPU is larger than the 126 MB L2 (no flush needed)",
                      "parallelism": f"reads sharded over {world} GPU(s), no collective",
                      "emitted_labels_per_row": emitted / (N * T), "reads_ok": ok_reads},
           "clocks": clocks, "gpu_launches": launches_per_step * steps, "roofline": roofline}

    # ---- end to end through the C ABI with pinned HOST buffers (copies inside the timed region);
    #      every rank streams its own shard, time = max over ranks ------------------------------------
    ne = min(w["e2e_reads"], N)
    hpost = torch.empty((ne, T, Cn), dtype=post.dtype).pin_memory()
    hpost.copy_(post[:ne])
    hnp = hpost.numpy()
    hcap = ne * stride
    hnp2 = henv = None
    if w["kind"] == "duplex":
        hpost2 = torch.empty((ne, w["T2"], Cn), dtype=torch.float32).pin_memory()
        hpost2.copy_(post2[:ne])
        henv_t = torch.from_numpy(band_envelope_np(ne, T, w["T2"], w["band"])).pin_memory()  # pinned like the posteriors
        henv = henv_t.numpy()
        hnp2 = hpost2.numpy()
        outd = None
    else:
        hl = torch.empty(hcap, dtype=torch.uint8).pin_memory()
        hq = torch.empty(hcap, dtype=torch.uint8).pin_memory()
        hp = torch.empty(hcap, dtype=torch.int32).pin_memory()
        ho = torch.empty(ne, dtype=torch.int64).pin_memory()
        hn = torch.empty(ne, dtype=torch.int32).pin_memory()
        hs = torch.empty(ne, dtype=torch.int32).pin_memory()
        outd = dict(labels=hl.numpy(), quals=hq.numpy(), path=hp.numpy(), offsets=ho.numpy(), lens=hn.numpy(),
                    status=hs.numpy())

    # the device-resident arrays of the first measurement are done with: the host-memory entry points
    # size their own device buffers from what is free, as they would in a caller's process
    torch.cuda.synchronize()
    del post, labels, path, offsets, lens, status, used, scratch, quals, post2, env, step
    torch.cuda.empty_cache()

    def e2e_step():
        if w["kind"] == "viterbi":
            return dec.viterbi_batch(hnp, bool(w["qstring"]), w["qscale"], w["qbias"], bool(w["collapse"]),
                                     qual_bytes=1, out=outd, packed=bool(w.get("e2e_packed", True)))
        if w["kind"] == "duplex":
            return dec.beam_search_duplex_batch(hnp, hnp2, henv, w["beam"], w["thr"], bool(w["collapse"]), w["exp_mode"])
        return dec.beam_search_batch(hnp, w["beam"], w["thr"], bool(w["collapse"]), out=outd)
    for _ in range(2 if w["kind"] != "viterbi" else 3):
        res_e = e2e_step()
    barrier()
    reps = w.get("e2e_reps", 5)
    t0 = time.perf_counter()
    for _ in range(reps):
        res_e = e2e_step()
    dt = (time.perf_counter() - t0) / reps
    if world > 1:
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
    h2d = ne * T * Cn * eb
    if w["kind"] == "viterbi":
        d2h = int(res_e["d2h_bytes"]) if "d2h_bytes" in res_e else hcap * (1 + 4 + (1 if w["qstring"] else 0)) + ne * 16
    elif w["kind"] == "duplex":
        h2d = ne * (T + w["T2"]) * Cn * 4 + ne * T * 16
        d2h = int(res_e["used"]) + ne * 16 + 8
    else:
        d2h = int(res_e["used"]) * 5 + ne * 16 + 8
    res["e2e"] = {"value": world * ne * T / dt, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                  "reads_per_step_per_gpu": ne, "ms_per_step": dt * 1e3,
                  "h2d_gbs_per_gpu": h2d / dt / 1e9, "d2h_gbs_per_gpu": d2h / dt / 1e9,
                  "note": "fctc_*_batch (host pointers, pinned memory); every rank streams its own shard, max over ranks"}
    if rank == 0 and world == 1:
        # ---- CPU baseline on a bounded sample of the same bytes (N = 1 only: under torchrun the ranks share the host
        #      cores, and rank 0 is pinned to its share of them; the reference arm gives the CPU figure there) --------
        import oracle
        oracle.build()
        nc = min(w["cpu_sample"], ne)
        threads = host_threads()
        sample = (hnp[:nc], hnp2[:nc], henv[:nc]) if w["kind"] == "duplex" else hnp[:nc]
        if f16:  # the CPU path only takes float32: it decodes the widened sample (widening outside its timed region)
            sample = sample.astype(np.float32)
        rate, dtc, passes = cpu_rate(w, sample, threads, min_seconds=w.get("cpu_seconds", 1.0))
        res["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
                               "sample": f"first {nc} reads of rank 0's shard ({nc * T} rows), {passes} pass(es), {dtc:.1f} s "
                                         f"of wall clock on {threads} threads, one read per worker thread; C++ restatement "
                                         "of the reference (Rust toolchain absent)"}
    barrier()
    del hpost, hnp, outd
    return res


def run_ours(args, names):
    import torch
    import torch.distributed as dist

    import fctc_b200

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", str(world)))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the decoder has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    binding = bind_to_gpu_numa_node(torch, local_rank, local_world)
    if world > 1:
        # stdout carries exactly one JSON line: while the communicator comes up (NCCL prints its version
        # banner and any NCCL_DEBUG output to stdout) file descriptor 1 points at stderr
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)
    dec = fctc_b200.Decoder(local_rank)
    ctx = dict(torch=torch, dist=dist, dec=dec, rank=rank, local_rank=local_rank, world=world, dev=dev)

    head = measure_workload(names[0], WORKLOADS[names[0]], args, args.steps, ctx)
    line = {"metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": head["warmup"], "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": head["dtype"], "data": "synthetic", "config": head["config"],
            "clocks": head["clocks"], "gpu_launches": head["gpu_launches"], "roofline": head["roofline"],
            "e2e": head["e2e"]}
    line["config"]["host_binding"] = binding
    if "cpu_baseline" in head:
        line["cpu_baseline"] = head["cpu_baseline"]
    if len(names) > 1:
        line["workloads"] = {}
        for nm in names[1:]:
            w = WORKLOADS[nm]
            line["workloads"][nm] = measure_workload(nm, w, args, max(1, min(args.steps, w.get("max_steps", 5))), ctx)
        line["gpu_launches"] += sum(v["gpu_launches"] for v in line["workloads"].values())
    if rank == 0:
        print(json.dumps(line))
    dec.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


ALL = ["viterbi", "viterbi_f16", "beam5", "beam32", "duplex", "duplex_cuda_exp"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=sorted(WORKLOADS) + ["all"], default="all",
                    help="all (default): the Viterbi headline plus the other BASELINE workloads under \"workloads\"")
    ap.add_argument("--reads", type=int, default=0, help="override reads per GPU (development only)")
    args = ap.parse_args()
    names = ALL if args.workload == "all" else [args.workload]
    if args.impl == "reference":
        run_reference(args, names)
    else:
        run_ours(args, names)


if __name__ == "__main__":
    main()
