#!/usr/bin/env python3
"""Full-size parity report: the BASELINE.json batches decoded on the GPU (through the C ABI) and by the CPU oracle on
the same bytes, compared read by read.  Written once per round on the GPU box; the result is committed as
profiles/PARITY_rNN.json.

  Viterbi   all 1,000,000 reads x 4000 x 5 (qstring, qscale 0.7, qbias 2.0): labels, path and quality characters
  beam 5    all 262,144 reads x 4000 x 5, threshold 0.1: labels and path
  beam 32   8,192 reads (default) of the 65,536 x 10,000 x 5 batch, threshold 0: labels and path
  duplex    2,048 pairs (default) of the 16,384-pair batch, band +-32, beam 5: labels, exp_mode 0 and 1

The inputs are the device generator's (the bytes bench.py decodes, seed 0xC7C00001), streamed in chunks: every
chunk is decoded on the device, copied to the host with its posteriors, decoded again by the oracle on all host
threads, and compared with numpy.  Reads that the device API flags FCTC_NODE_POOL_OVERFLOW are decoded again
through the host entry point (which retries with the hard bound), as a caller would.

  python tools/parity_full.py [--out gpurun_out/PARITY.json] [--viterbi N] [--beam5 N] [--beam32 N] [--duplex N]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "fast-ctc-decode_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402

import fctc_b200  # noqa: E402
import oracle  # noqa: E402

SEED = 0xC7C00001


def gather_pool(arr, offsets, lens):
    """Concatenation of arr[offsets[n] : offsets[n] + lens[n]] over n."""
    lens = lens.astype(np.int64)
    total = int(lens.sum())
    if total == 0:
        return arr[:0]
    starts = np.repeat(offsets.astype(np.int64) - np.concatenate([[0], np.cumsum(lens)[:-1]]), lens)
    return arr[starts + np.arange(total, dtype=np.int64)]


def gather_rows(mat, lens):
    """Concatenation of mat[n, :lens[n]] over n (oracle layout [N, T])."""
    T = mat.shape[1]
    mask = np.arange(T)[None, :] < lens[:, None]
    return mat[mask]


def first_bad_reads(eq_flat, lens, limit=8):
    """Indices of the reads that hold a False of eq_flat (concatenated per-read entries)."""
    if eq_flat.all():
        return []
    owner = np.repeat(np.arange(lens.size), lens.astype(np.int64))
    return np.unique(owner[~eq_flat])[:limit].tolist()


def viterbi_report(dec, n_reads, chunk=8192):
    T, C, stride = 4000, 5, 4000
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    post = torch.empty((chunk, T, C), dtype=torch.float32, device=dev)
    cap = chunk * stride
    labels = torch.empty(cap, dtype=torch.uint8, device=dev)
    quals = torch.empty(cap, dtype=torch.uint8, device=dev)
    path = torch.empty(cap, dtype=torch.int32, device=dev)
    offsets = torch.empty(chunk, dtype=torch.int64, device=dev)
    lens = torch.empty(chunk, dtype=torch.int32, device=dev)
    status = torch.empty(chunk, dtype=torch.int32, device=dev)
    rep = dict(reads=0, identical=0, labels_compared=0, differing_reads=[], t_oracle_s=0.0)
    for lo in range(0, n_reads, chunk):
        cnt = min(chunk, n_reads - lo)
        dec.synth_posteriors_device(post.data_ptr(), cnt, T, C, 0, SEED, lo, st)
        dec.viterbi_batch_device(post.data_ptr(), cnt, T, C, 1, 0.7, 2.0, 1, labels.data_ptr(), quals.data_ptr(), 1,
                                 path.data_ptr(), cap, offsets.data_ptr(), lens.data_ptr(), status.data_ptr(), st)
        torch.cuda.synchronize()
        hp = post[:cnt].cpu().numpy()
        t0 = time.perf_counter()
        ol, oq, op, olen, ost = oracle.viterbi_batch(hp, True, 0.7, 2.0, True)
        rep["t_oracle_s"] += time.perf_counter() - t0
        glen = lens[:cnt].cpu().numpy().astype(np.int64)
        gst = status[:cnt].cpu().numpy()
        gl = labels.view(chunk, stride)[:cnt].cpu().numpy()
        gq = quals.view(chunk, stride)[:cnt].cpu().numpy()
        gp = path.view(chunk, stride)[:cnt].cpu().numpy()
        ok = (glen == olen) & (gst == 0) & (ost == 0)
        same_len = glen == olen
        m = np.minimum(glen, olen)
        eq = (gather_rows(gl, m) == gather_rows(ol, m)) & (gather_rows(gp, m) == gather_rows(op, m)) & \
             (gather_rows(gq, m).astype(np.uint32) == gather_rows(oq, m))
        bad = set(first_bad_reads(eq, m, 1 << 30)) | set(np.flatnonzero(~ok).tolist())
        rep["reads"] += cnt
        rep["identical"] += cnt - len(bad)
        rep["labels_compared"] += int(m.sum())
        for n in sorted(bad)[:8]:
            if len(rep["differing_reads"]) < 32:
                rep["differing_reads"].append(dict(read=lo + n, gpu_len=int(glen[n]), oracle_len=int(olen[n]),
                                                   gpu_status=int(gst[n]), same_len=bool(same_len[n])))
    rep["what"] = "labels, emission rows and quality characters of every read, bit for bit"
    return rep


def beam_report(dec, name, n_total, n_reads, T, beam, thr, chunk):
    C = 5
    stride = (T + 31) // 32 * 32
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    post = torch.empty((chunk, T, C), dtype=torch.float32, device=dev)
    cap = chunk * stride
    labels = torch.empty(cap, dtype=torch.uint8, device=dev)
    path = torch.empty(cap, dtype=torch.int32, device=dev)
    offsets = torch.empty(chunk, dtype=torch.int64, device=dev)
    lens = torch.empty(chunk, dtype=torch.int32, device=dev)
    status = torch.empty(chunk, dtype=torch.int32, device=dev)
    used = torch.zeros(1, dtype=torch.int64, device=dev)
    free_b, _ = torch.cuda.mem_get_info(0)
    nbytes = min(dec.beam_scratch_bytes(chunk, T, C, beam, 0), int(free_b * 0.8) // 256 * 256)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    # reads spread over the whole batch when only part of it is compared
    step = max(1, n_total // n_reads)
    rep = dict(reads=0, identical=0, labels_compared=0, pool_overflow_retried=0, differing_reads=[], t_oracle_s=0.0,
               sampling=f"reads {0}, {step}, {2 * step}, ... of the {n_total}-read batch" if step > 1 else "every read")
    for lo in range(0, n_reads, chunk):
        cnt = min(chunk, n_reads - lo)
        if step == 1:
            dec.synth_posteriors_device(post.data_ptr(), cnt, T, C, 0, SEED, lo, st)
        else:
            for j in range(cnt):  # one read at a time: the generator is keyed by the read's index in the batch
                dec.synth_posteriors_device(post[j].data_ptr(), 1, T, C, 0, SEED, (lo + j) * step, st)
        dec.beam_search_batch_device(post.data_ptr(), cnt, T, C, beam, thr, True, labels.data_ptr(), path.data_ptr(), cap,
                                     offsets.data_ptr(), lens.data_ptr(), status.data_ptr(), used.data_ptr(),
                                     scratch.data_ptr(), nbytes, 0, st)
        torch.cuda.synchronize()
        hp = post[:cnt].cpu().numpy()
        t0 = time.perf_counter()
        ol, op, olen, ost, _ = oracle.beam_search_batch(hp, beam, thr, True)
        rep["t_oracle_s"] += time.perf_counter() - t0
        glen = lens[:cnt].cpu().numpy().astype(np.int64)
        gst = status[:cnt].cpu().numpy().copy()
        goff = offsets[:cnt].cpu().numpy()
        gl, gp = labels.cpu().numpy(), path.cpu().numpy()
        retried = {}
        for n in np.flatnonzero(gst == 4):  # trie pool too small at the default plan: the host entry point retries
            s1, l1, p1 = dec.beam_search(hp[n], beam, thr, True)
            retried[int(n)] = (s1, l1, p1)
            rep["pool_overflow_retried"] += 1
        m = np.where(gst == 0, np.minimum(glen, olen), 0)
        eq = (gather_pool(gl, goff, m) == gather_rows(ol, m)) & (gather_pool(gp, goff, m) == gather_rows(op, m))
        bad = set(first_bad_reads(eq, m, 1 << 30))
        bad |= set(np.flatnonzero((gst == 0) & ((glen != olen) | (ost != 0))).tolist())
        bad |= set(np.flatnonzero((gst != 0) & (gst != 4) & (gst != ost)).tolist())
        for n, (s1, l1, p1) in retried.items():
            if not (s1 == ost[n] and l1.tolist() == ol[n, :olen[n]].tolist() and p1.tolist() == op[n, :olen[n]].tolist()):
                bad.add(n)
        rep["reads"] += cnt
        rep["identical"] += cnt - len(bad)
        rep["labels_compared"] += int(m.sum())
        for n in sorted(bad)[:8]:
            if len(rep["differing_reads"]) < 32:
                rep["differing_reads"].append(dict(read=(lo + n) * step, gpu_len=int(glen[n]), oracle_len=int(olen[n]),
                                                   gpu_status=int(gst[n]), oracle_status=int(ost[n])))
    rep["what"] = f"{name}: labels and first-proposal rows of every compared read, exact"
    return rep


def duplex_report(dec, n_total, n_pairs, mode, chunk=1024):
    T, C, band, beam = 4000, 5, 32, 5
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    c = np.arange(T)
    env1 = np.stack([np.maximum(0, c - band), np.minimum(T, c + band + 1)], axis=1).astype(np.int64)
    step = max(1, n_total // n_pairs)
    rep = dict(pairs=0, identical=0, labels_compared=0, differing_pairs=[], t_oracle_s=0.0, exp_mode=mode,
               sampling=f"pairs 0, {step}, {2 * step}, ... of the {n_total}-pair batch" if step > 1 else "every pair")
    a = torch.empty((chunk, T, C), dtype=torch.float32, device=dev)
    b = torch.empty((chunk, T, C), dtype=torch.float32, device=dev)
    for lo in range(0, n_pairs, chunk):
        cnt = min(chunk, n_pairs - lo)
        for j in range(cnt):
            dec.synth_posteriors_device(a[j].data_ptr(), 1, T, C, 1, SEED, (lo + j) * step, st)
            dec.synth_posteriors_device(b[j].data_ptr(), 1, T, C, 2, SEED, (lo + j) * step, st)
        torch.cuda.synchronize()
        ha, hb = a[:cnt].cpu().numpy(), b[:cnt].cpu().numpy()
        env = np.ascontiguousarray(np.broadcast_to(env1, (cnt, T, 2)))
        res = dec.beam_search_duplex_batch(ha, hb, env, beam, 0.0, True, exp_mode=mode)
        t0 = time.perf_counter()
        ol, olen, ost = oracle.beam_search_duplex_batch(ha, hb, env.astype(np.uint64), beam, 0.0, True, mode)
        rep["t_oracle_s"] += time.perf_counter() - t0
        glen = res["lens"].astype(np.int64)
        m = np.where(res["status"] == 0, np.minimum(glen, olen), 0)
        eq = gather_pool(res["labels"], res["offsets"], m) == gather_rows(ol, m)
        bad = set(first_bad_reads(eq, m, 1 << 30)) | set(np.flatnonzero((glen != olen) | (res["status"] != ost)).tolist())
        rep["pairs"] += cnt
        rep["identical"] += cnt - len(bad)
        rep["labels_compared"] += int(m.sum())
        for n in sorted(bad)[:8]:
            if len(rep["differing_pairs"]) < 32:
                rep["differing_pairs"].append(dict(pair=(lo + n) * step, gpu_len=int(glen[n]), oracle_len=int(olen[n]),
                                                   gpu_status=int(res["status"][n]), oracle_status=int(ost[n])))
    rep["what"] = "labels of every compared pair, exact (exp_mode 0: glibc expf/log1pf on both sides; 1: max-plus)"
    return rep


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "PARITY.json"))
    ap.add_argument("--viterbi", type=int, default=1_000_000)
    ap.add_argument("--beam5", type=int, default=262_144)
    ap.add_argument("--beam32", type=int, default=8192)
    ap.add_argument("--duplex", type=int, default=2048)
    args = ap.parse_args()
    oracle.build()
    dec = fctc_b200.Decoder(0)
    out = {"inputs": "device generator (fctc_synth_posteriors_device), seed 0xC7C00001: the bytes bench.py decodes",
           "oracle": "oracle/ctc_oracle.cpp on %d host threads" % oracle.hardware_threads(),
           "gpu": torch.cuda.get_device_name(0)}
    t0 = time.time()
    if args.viterbi:
        out["viterbi"] = viterbi_report(dec, args.viterbi)
        print("viterbi", out["viterbi"]["identical"], "/", out["viterbi"]["reads"], flush=True)
    if args.beam5:
        out["beam5"] = beam_report(dec, "beam 5, threshold 0.1", 262_144, args.beam5, 4000, 5, 0.1, 16384)
        print("beam5", out["beam5"]["identical"], "/", out["beam5"]["reads"], flush=True)
    if args.beam32:
        out["beam32"] = beam_report(dec, "beam 32, threshold 0", 65_536, args.beam32, 10_000, 32, 0.0, 2048)
        print("beam32", out["beam32"]["identical"], "/", out["beam32"]["reads"], flush=True)
    if args.duplex:
        for mode in (0, 1):
            key = "duplex_exp_mode_%d" % mode
            out[key] = duplex_report(dec, 16_384, args.duplex, mode)
            print(key, out[key]["identical"], "/", out[key]["pairs"], flush=True)
    out["wall_s"] = time.time() - t0
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    with open(args.out, "w") as f:
        json.dump(out, f, indent=1)
    dec.close()


if __name__ == "__main__":
    main()
