"""Generates the golden fixtures in this directory from the reference's own code compiled in place
(oracle/_ref/libfsq_ref.so, built by `make -C oracle ref` from /root/reference; see oracle/ref_harness.cc).

    python tests/golden/make_golden.py

The reference ships no tests or fixtures for this path; these files are what pins the oracle
(and through it the CUDA path) on machines where /root/reference does not exist.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import oraclelib  # noqa: E402
from __graft_entry__ import load_synth  # noqa: E402

synth = load_synth()
REF = oraclelib.ref()


def case(name, rows, L, dump_cols=40):
    out = {"rows": rows, "L": np.uint64(L)}
    with REF.open(rows) as r:
        out["comp"] = r.alphabet()
        nc = min(dump_cols, rows.shape[1])
        a, d, off, hv, hc = r.scan_dump(nc)
        out.update(scan_a=a, scan_d=d, hist_off=off, hist_v=hv, hist_c=hc)
        status = r.segment(L, True)
        out["status"] = np.int32(status)
        if status == 1:
            row, ln = r.short_path()
            out.update(sp_row=row, sp_len=ln)
        else:
            lb, rb, mx, sz = r.dp()
            out.update(dp_lb=lb, dp_rb=rb, dp_max=mx, dp_size=sz)
            lb, rb, mx, sz = r.traceback()
            out.update(tb_lb=lb, tb_rb=rb, tb_max=mx, tb_size=sz, K=np.uint32(r.K()))
        if status == 0:
            lb, rb, mx, sz = r.reduced()
            a, d = r.reduced_states()
            out.update(red_lb=lb, red_rb=rb, red_size=sz, red_a=a, red_d=d)
            offs, ids, cps = [], [], []
            for s in range(len(lb)):
                o, i, c = r.segment_texts(s)
                offs.append(o); ids.append(i); cps.append(c)
            out["text_off"] = np.concatenate(offs)
            out["text_ids"] = np.concatenate(ids)
            out["text_copied_from"] = np.concatenate(cps)
            if len(lb) > 1:
                out["weights01"] = r.ref_pair_weights(0)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, rows.shape, "L", L, "status", int(out["status"]), "K", int(out.get("K", 0)), "segments", len(out.get("red_lb", [])))


def main():
    kat = np.frombuffer(b"ACGTACGTAC" b"ACGTACGAAC" b"ACCTACGTAC" b"ACGTACGTAC" b"TCGTACGAAC", np.uint8).reshape(5, 10).copy()
    case("kat5x10_L2", kat, 2)
    case("kat5x10_short", kat, 6)
    case("synth16x400_L5", synth.synth_rows(16, 400, 1, 3), 5)
    case("synth48x1500_L10_gaps", synth.synth_rows(48, 1500, 7, 6, gaps=True), 10)
    case("synth64x5000_L10", synth.synth_rows(64, 5000, 1, 8), 10)
    case("synth200x700_L64", synth.synth_rows(200, 700, 5, 8, gaps=True, recomb_block=100), 64)
    rng = np.random.default_rng(0)
    case("binary12x80_L3", rng.choice(np.frombuffer(b"AC", np.uint8), size=(12, 80), p=[0.8, 0.2]), 3)
    case("random6x40_notreducible", rng.choice(np.frombuffer(b"ACGT", np.uint8), size=(6, 40)), 3)

    # An alignment on which founder_sequences::rmq's behaviour shows in the DP table: with an exact leftmost range-min in
    # its place, table (lb of hundreds of entries), traceback and reduced segmentation all come out different (found by
    # random search over low-diversity binary alignments, tests/golden/find_rmq_dp_quirk.py; few rows and rare changes
    # make the ranges between consecutive divergence values span many 64-entry blocks).
    case("rmq_dp_quirk_4x1599_L4", np.load(os.path.join(HERE, "rmq_dp_quirk_rows.npy")), 4)

    # rmq.hh, compiled as is: values that make the quirks visible
    vals = rng.integers(0, 1000, size=4096).astype(np.uint32)
    beg = rng.integers(0, 4095, size=3000).astype(np.uint64)
    end = np.minimum(4096, beg + rng.integers(1, 2500, size=3000).astype(np.uint64)).astype(np.uint64)
    ans = REF.rmq_queries(vals, beg, end)
    vals2 = rng.integers(0, 4, size=4096).astype(np.uint32)
    ans2 = REF.rmq_queries(vals2, beg, end)
    np.savez_compressed(os.path.join(HERE, "rmq_quirks.npz"), vals=vals, beg=beg, end=end, ans=ans, vals2=vals2, ans2=ans2)
    true_min = np.array([vals[b:e].min() for b, e in zip(beg, end)])
    print("rmq: %d of %d answers are not a minimum" % (int((vals[ans] != true_min).sum()), len(beg)))


if __name__ == "__main__":
    main()
