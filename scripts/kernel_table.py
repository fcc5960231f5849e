#!/usr/bin/env python
"""Per-kernel roofline table: every C-ABI entry point of the env path timed alone with CUDA events (after
warm-up, inputs larger than L2 or L2 flushed), algorithmic bytes per env from DESIGN.md section 3, GB/s and
fraction of the measured HBM peak.  One JSON line per (kernel, d)."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import toric_rl_decoder_b200 as pkg  # noqa: E402


def peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0


def timeit(fn, reps, flush):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(reps):
        flush.fill_(1)                       # 512 MB write: evicts the 126 MB L2
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ds", default="5,7,9,15,25")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--scale", type=int, default=1, help="multiply the env counts (asymptotic bandwidth)")
    a = ap.parse_args()
    pk = peak()
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    for d in [int(x) for x in a.ds.split(",")]:
        n = a.scale * ((1 << 20) if d <= 9 else ((1 << 18) if d <= 16 else (1 << 16)))
        eng = pkg.ToricEngine(d, n, seed=0, p_error=0.1)
        eng.reset()
        eng_lo = pkg.ToricEngine(d, n, seed=1, p_error=0.0005)
        eng_mid = pkg.ToricEngine(d, n, seed=1, p_error=0.005)
        acts = torch.stack([torch.randint(0, 2, (n,)), torch.randint(0, d, (n,)), torch.randint(0, d, (n,)),
                            torch.randint(1, 4, (n,))], 1).to(torch.int32).cuda()
        eng.actions.copy_(acts)
        u8 = eng.get_state()
        W = 4   # bytes per packed row word
        rows = [
            ("tc_syndrome (K2)", lambda: eng.syndrome(), 4 * d * W + 2 * d * W),
            ("tc_step (K3)", lambda: eng.lib.tc_step(pkg.engine._ptr(eng.qubits), pkg.engine._ptr(eng.syn),
                                                     pkg.engine._ptr(eng.actions), pkg.engine._ptr(eng.reward),
                                                     pkg.engine._ptr(eng.terminal), pkg.engine._ptr(eng.status), d, n,
                                                     pkg.engine._stream()), 2 * 6 * d * W + 16 + 5),
            ("tc_ground_state (K5)", lambda: eng.ground_state(), 4 * d * W + 1),
            ("tc_reset p=0.1 (K1)", lambda: eng.reset(), 6 * d * W + 4 + 8 + 1),
            ("tc_reset p=0.005 (K1)", lambda: eng_mid.reset(), 6 * d * W + 4 + 8 + 1),
            ("tc_reset p=0.0005 (K1)", lambda: eng_lo.reset(), 6 * d * W + 4 + 8 + 1),
            ("tc_reset fixed weight 3 (K1)", lambda: eng_lo.reset(mode=1, n_fixed=3), 6 * d * W + 4 + 8 + 1),
            ("tc_persp_count + scan", lambda: eng.perspectives_count(), 2 * d * W + 4 + 4),
            ("tc_unpack_state", lambda: eng.get_state(), 2 * d * W + 2 * d * d),
            ("tc_pack_state", lambda: eng.set_state(u8), 2 * d * W + 2 * d * d),
            ("tc_actor_pre (fused step+reset+count)", lambda: eng.actor_pre(), 2 * 6 * d * W + 16 + 4 + 1 + 1 + 8 + 8 + 8),
        ]
        for name, fn, bpe in rows:
            ms = timeit(fn, a.reps, flush)
            gbs = bpe * n / (ms * 1e-3) / 1e9
            print(json.dumps({"kernel": name, "d": d, "envs": n, "ms": round(ms, 5), "alg_bytes_per_env": bpe,
                              "GBps": round(gbs, 1), "frac_of_measured_peak": round(gbs / pk, 3),
                              "envs_per_s": round(n / (ms * 1e-3))}), flush=True)
        del eng


if __name__ == "__main__":
    main()
