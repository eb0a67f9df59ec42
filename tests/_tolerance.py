"""The parity rule of the tests, in one place, plus the bookkeeping VERDICT r1 asked for.

north_star: 1e-4 relative / 1e-5 absolute in fp32 for outputs and all gradients.  Parameter gradients
are fp32 sums of up to N*H*W products whose magnitude grows with the batch, so for them the tests scale
the ABSOLUTE term by max(1, max|ref|) of the tensor (`scale_atol=True`).  That is a
concession for near-zero entries of a tensor whose large entries pass the relative term, and its size
is measured, not assumed: every call also counts the elements that fail the UN-scaled 1e-4 / 1e-5 rule
and the largest error in units of (1e-5 + 1e-4 |ref|); conftest.py prints the totals at the end of the
run and writes them to gpurun_out/strict_tolerance.json.
"""
import numpy as np

RTOL, ATOL = 1e-4, 1e-5
REPORT = {}          # tensor name -> [calls, elements, strict failures, worst strict ratio, worst max|ref|]


def _record(name, ref, err):
    strict_lim = ATOL + RTOL * np.abs(ref)
    ratio = err / strict_lim
    n_bad = int((ratio > 1.0).sum())
    key = name.split(" ")[0]
    r = REPORT.setdefault(key, [0, 0, 0, 0.0, 0.0])
    r[0] += 1; r[1] += int(err.size); r[2] += n_bad
    if err.size:
        worst = float(ratio.max())
        if worst > r[3]:
            r[3], r[4] = worst, float(np.abs(ref).max())
    return n_bad


def close(got, ref, name, scale_atol=False, rtol=RTOL, atol=ATOL):
    """Asserts |got - ref| <= atol' + rtol |ref| with atol' = atol * max(1, max|ref|) if scale_atol.
    Returns the number of elements outside the un-scaled rule (0 whenever scale_atol is False)."""
    got = np.asarray(got, np.float64)
    ref = np.asarray(ref, np.float64)
    assert got.shape == ref.shape, "%s: shape %s vs %s" % (name, got.shape, ref.shape)
    a = atol * (max(1.0, float(np.abs(ref).max())) if (scale_atol and ref.size) else 1.0)
    err = np.abs(got - ref)
    n_strict = _record(name, ref, err)
    bad = err > a + rtol * np.abs(ref)
    assert not bad.any(), "%s: %d/%d outside tolerance (%d outside the un-scaled rule), max abs err %.3e (max|ref| %.3e)" % (
        name, int(bad.sum()), bad.size, n_strict, float(err.max()), float(np.abs(ref).max()))
    return n_strict


def summary_lines():
    if not REPORT:
        return []
    out = ["un-scaled 1e-4 rel / 1e-5 abs rule, per tensor kind: calls, elements, elements outside, worst error in "
           "units of the un-scaled limit (max|ref| of that tensor)"]
    for k in sorted(REPORT):
        c, n, b, w, m = REPORT[k]
        out.append("  %-28s %5d calls %12d elements %9d outside (%.2e)  worst %.2fx  (max|ref| %.3g)" % (
            k, c, n, b, b / max(1, n), w, m))
    return out
