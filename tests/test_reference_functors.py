"""Pins the residual models to the REFERENCE'S OWN SOURCE: tests/golden/reference_functor_vectors.npz holds residuals and Jacobian blocks
computed by the reference's cost functors themselves (BundleAdjuster::ReprojectCost, src/adjuster/BundleAdjuster.h:33-69, and
SnavelyReprojectionError, src/adjuster/SnavelyReprojectionError.h) compiled here by oracle/ref_build.sh and differentiated with Jets as
Ceres' AutoDiffCostFunction does (generator: tests/golden/make_reference_vectors.py).  CPU: the oracle against those vectors, and — where
/root/reference exists — the functor library against the committed file.  GPU: the CUDA engine against the same vectors.
What this does NOT pin: ceres::AngleAxisRotatePoint (Ceres is absent; oracle/shim restates it) and the trust-region loop."""
import os

import numpy as np
import pytest

import reffunctor as rf

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_functor_vectors.npz")
KEYS = ("pose", "intr", "cam_group", "cam_fixed", "pt", "obs_cam", "obs_pt", "obs_uv")


def cases(ba):
    d = np.load(GOLD)
    pin = ba.Problem(**{k: d[f"pin_{k}"] for k in KEYS})
    sna = ba.Problem(**{k: d[f"sna_{k}"] for k in KEYS})
    yield "pinhole rule 0", pin, dict(model=0, taylor_rule=0), {k: d[f"pin_{k}_rule0"] for k in ("r", "Ji", "Jc", "Jp")}
    yield "pinhole rule 1", pin, dict(model=0, taylor_rule=1), {k: d[f"pin_{k}_rule1"] for k in ("r", "Ji", "Jc", "Jp")}
    yield "snavely", sna, dict(model=1, taylor_rule=0), {k: d[f"sna_{k}"] for k in ("r", "Ji", "Jc", "Jp")}


def check(ba, make_handle, rtol, only=None):
    n = 0
    for name, P, opts, ref in cases(ba):
        if only and name not in only:
            continue
        with make_handle(intrinsics_free=1, robust_kind=0, **opts) as h:
            h.set_problem(P)
            e = h.evaluate()
        free = (P.cam_fixed == 0)[P.obs_cam][:, None, None]     # a constant pose has no Jacobian columns (SetParameterBlockConstant, :105)
        for k, want in (("r", ref["r"]), ("Jc", ref["Jc"] * free), ("Jp", ref["Jp"]), ("Ji", ref["Ji"])):
            scale = np.abs(want).max()
            assert np.isfinite(want).all() and scale > 0
            assert np.abs(e[k] - want).max() <= rtol * scale, (name, k, np.abs(e[k] - want).max() / scale)
        n += 1
    assert n > 0


def test_oracle_matches_the_reference_functors(ba, oracle):
    check(ba, lambda **kw: oracle.create(oracle.default_options(**kw)), rtol=1e-13)


@pytest.mark.skipif(not rf.available(), reason="oracle/_ref/libref_cost.so is built only where /root/reference exists")
def test_committed_vectors_come_from_the_reference_build(ba):
    for name, P, opts, ref in cases(ba):
        r, Ji, Jc, Jp = rf.evaluate_problem(opts["model"], P, opts["taylor_rule"])
        for got, want in ((r, ref["r"]), (Ji, ref["Ji"]), (Jc, ref["Jc"]), (Jp, ref["Jp"])):
            assert np.array_equal(got, want), name


@pytest.mark.gpu
def test_engine_matches_the_reference_functors(ba):
    """Analytic device Jacobians (K0 + K1) against the reference functor's autodiff: the same 1e-10-of-scale bar as engine vs oracle, with margin."""
    check(ba, lambda **kw: ba.Engine(**kw), rtol=1e-9, only=("pinhole rule 0", "snavely"))


@pytest.mark.gpu
def test_engine_matches_the_reference_functors_taylor_rule_1(ba):
    """taylor_rule = 1 (first-order branch only for an exactly zero rotation): a rotation of ~1e-9 rad then goes through Rodrigues' formula,
    where the functor's Jet arithmetic forms (1 - cos t)/t^2 from a cancelled numerator while K0 uses the power series: the two agree to the
    size of that term (1e-9 of scale), hence the looser bar of this one case."""
    check(ba, lambda **kw: ba.Engine(**kw), rtol=1e-7, only=("pinhole rule 1",))
