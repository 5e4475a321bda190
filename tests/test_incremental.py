"""Incremental use (SequentialActuator.h:138-229 runs a full BA after every added image): ba_append grows the problem
held by a handle — state kept, new ids appended, observations merged into point order.  CPU: the oracle's implementation and the
shared merge (include/ba_merge.hpp).  GPU: the engine against a fresh handle on the merged problem (bit-exact) and against the oracle."""
import numpy as np
import pytest


def staged(ba, name, cuts, seed=0):
    """Re-index a synthetic scene so that it can be grown camera range by camera range: points are numbered by the stage in which
    they are first seen.  Returns the full problem and [(Nc_s, Np_s)] per stage."""
    P = ba.synth(name, seed=seed)
    Nc, _, Np, No = P.sizes
    stage_of_cam = np.searchsorted(np.asarray(cuts), np.arange(Nc), side="right")
    first = np.full(Np, len(cuts), np.int64)
    np.minimum.at(first, P.obs_pt, stage_of_cam[P.obs_cam])
    order = np.lexsort((np.arange(Np), first))                 # points by (first stage, old id); never-observed points last
    new_id = np.empty(Np, np.int64); new_id[order] = np.arange(Np)
    opt = new_id[P.obs_pt]
    perm = np.argsort(opt, kind="stable")
    Q = ba.Problem(pose=P.pose, intr=P.intr, cam_group=P.cam_group, cam_fixed=P.cam_fixed, pt=P.pt.reshape(-1, 3)[order], obs_cam=P.obs_cam[perm],
                   obs_pt=opt[perm].astype(np.int32), obs_uv=P.obs_uv.reshape(-1, 2)[perm])
    sizes = [(int(c), int((first <= s).sum())) for s, c in enumerate(cuts)]
    return Q, sizes


def sub_problem(ba, Q, nc, npts, pose=None, pt=None):
    keep = Q.obs_cam < nc
    assert (Q.obs_pt[keep] < npts).all()
    pose = Q.pose.reshape(-1, 6) if pose is None else pose
    pt = Q.pt.reshape(-1, 3) if pt is None else pt
    return ba.Problem(pose=pose[:nc], intr=Q.intr, cam_group=Q.cam_group[:nc], cam_fixed=Q.cam_fixed[:nc], pt=pt[:npts], obs_cam=Q.obs_cam[keep],
                      obs_pt=Q.obs_pt[keep], obs_uv=Q.obs_uv.reshape(-1, 2)[keep])


def delta(Q, nc0, np0, nc1, np1, shuffle=None):
    new = (Q.obs_cam >= nc0) & (Q.obs_cam < nc1)
    idx = np.flatnonzero(new)
    if shuffle is not None:
        idx = shuffle.permutation(idx)
    return dict(pose=Q.pose.reshape(-1, 6)[nc0:nc1], cam_group=Q.cam_group[nc0:nc1], cam_fixed=Q.cam_fixed[nc0:nc1], pt=Q.pt.reshape(-1, 3)[np0:np1],
                obs_cam=Q.obs_cam[idx], obs_pt=Q.obs_pt[idx], obs_uv=Q.obs_uv.reshape(-1, 2)[idx])


def check_merge_order(make, ba):
    Q, sizes = staged(ba, "sequential30", [10, 20, 30])
    (nc0, np0), (nc1, np1) = sizes[0], sizes[1]
    rng = np.random.default_rng(1)
    with make() as h:
        h.set_problem(sub_problem(ba, Q, nc0, np0))
        d = delta(Q, nc0, np0, nc1, np1, shuffle=rng)
        h.append(**d)
        assert h.sizes == (nc1, 1, np1, int((Q.obs_cam < nc1).sum()))
        oc, op, uv = h.observations()
        pose, intr, pt = h.get_state()
    assert np.all(np.diff(op) >= 0)
    # within a point: the existing observations (cameras < nc0, original order) first, then the new ones in the order given
    base = sub_problem(ba, Q, nc0, np0)
    for p in (0, 5, np0 - 1, np0, np1 - 1):
        got = [(int(c), tuple(x)) for c, x in zip(oc[op == p], uv[op == p])]
        old = [(int(c), tuple(x)) for c, x in zip(base.obs_cam[base.obs_pt == p], base.obs_uv.reshape(-1, 2)[base.obs_pt == p])]
        new = [(int(c), tuple(x)) for c, x in zip(d["obs_cam"][d["obs_pt"] == p], d["obs_uv"][d["obs_pt"] == p])]
        assert got == old + new
    assert np.array_equal(pose, Q.pose.reshape(-1, 6)[:nc1]) and np.array_equal(pt, Q.pt.reshape(-1, 3)[:np1]) and np.array_equal(intr, Q.intr.reshape(-1, 4))


def check_append_equals_fresh_problem(make, ba, exact):
    """solve(stage 0); append(stage 1); solve  ==  a fresh handle given the merged problem with the same starting state."""
    Q, sizes = staged(ba, "sequential30", [10, 20, 30])
    (nc0, np0), (nc1, np1) = sizes[0], sizes[1]
    with make() as h:
        h.set_problem(sub_problem(ba, Q, nc0, np0))
        s0 = h.solve()
        pose0, _, pt0 = h.get_state()
        h.append(**delta(Q, nc0, np0, nc1, np1))
        pose_m, intr_m, pt_m = h.get_state()
        assert np.array_equal(pose_m[:nc0], pose0) and np.array_equal(pt_m[:np0], pt0)                      # the solved state is kept
        assert np.array_equal(pose_m[nc0:], Q.pose.reshape(-1, 6)[nc0:nc1]) and np.array_equal(pt_m[np0:], Q.pt.reshape(-1, 3)[np0:np1])
        s1 = h.solve()
        t1 = h.trajectory()
        x1 = h.get_state()
        h.reset_state()                                                                                      # restores the MERGED starting state
        assert all(np.array_equal(a, b) for a, b in zip(h.get_state(), (pose_m, intr_m, pt_m)))
    assert s0.usable and s1.usable and s1.initial_cost > s0.final_cost
    M = sub_problem(ba, Q, nc1, np1, pose=pose_m, pt=pt_m)
    M = ba.Problem(pose=M.pose, intr=intr_m, cam_group=M.cam_group, cam_fixed=M.cam_fixed, pt=M.pt, obs_cam=M.obs_cam, obs_pt=M.obs_pt, obs_uv=M.obs_uv)
    with make() as f:
        f.set_problem(M)
        f.solve()
        t2 = f.trajectory()
        x2 = f.get_state()
    assert len(t1) == len(t2)
    for a, b in zip(t1, t2):
        assert a.step_is_successful == b.step_is_successful
        assert a.cost == b.cost if exact else abs(a.cost - b.cost) <= 1e-9 * abs(b.cost)
    for a, b in zip(x1, x2):
        assert np.array_equal(a, b) if exact else np.allclose(a, b, rtol=1e-7, atol=1e-9)


def check_validation(make, ba):
    P = ba.synth("tiny")
    with make() as h:
        with pytest.raises(ba.BAError):
            h.append(pt=np.zeros((1, 3)))                                                  # no problem set
        h.set_problem(P)
        Nc, _, Np, _ = P.sizes
        for bad in (dict(obs_cam=[Nc], obs_pt=[0], obs_uv=[[0, 0]]), dict(obs_cam=[0], obs_pt=[Np], obs_uv=[[0, 0]]),
                    dict(pose=np.zeros((1, 6)), cam_group=[1], cam_fixed=[0])):
            with pytest.raises(ba.BAError, match="ba_append"):
                h.append(**bad)
        h.append()                                                                          # an empty delta is fine
        assert h.sizes == P.sizes
        h.append(pt=np.ones((2, 3)), obs_cam=[0, 1], obs_pt=[Np + 1, Np + 1], obs_uv=[[1, 2], [3, 4]])   # point Np stays unobserved
        assert h.sizes == (Nc, 1, Np + 2, P.sizes[3] + 2)
        assert h.solve().usable


def oracle_maker(oracle, **kw):
    return lambda: oracle.create(oracle.default_options(num_threads=4, max_num_iterations=8, **kw))


def test_oracle_append_merge_order(ba, oracle):
    check_merge_order(oracle_maker(oracle), ba)


def test_oracle_append_equals_fresh_problem(ba, oracle):
    check_append_equals_fresh_problem(oracle_maker(oracle), ba, exact=True)


def test_oracle_append_validation(ba, oracle):
    check_validation(oracle_maker(oracle), ba)


@pytest.mark.gpu
def test_engine_append_merge_order(ba):
    check_merge_order(lambda: ba.Engine(max_num_iterations=8), ba)


@pytest.mark.gpu
@pytest.mark.parametrize("solver", [0, 1])
def test_engine_append_equals_fresh_problem(ba, solver):
    check_append_equals_fresh_problem(lambda: ba.Engine(max_num_iterations=8, linear_solver=solver), ba, exact=True)


@pytest.mark.gpu
def test_engine_append_validation(ba):
    check_validation(lambda: ba.Engine(max_num_iterations=8), ba)


@pytest.mark.gpu
def test_engine_growing_map_matches_oracle(ba, oracle):
    """The SequentialActuator pattern: BA after every batch of new images, engine and oracle side by side."""
    Q, sizes = staged(ba, "sequential30", [6, 12, 18, 24, 30], seed=3)
    with ba.Engine(max_num_iterations=6) as e, oracle.create(oracle.default_options(num_threads=4, max_num_iterations=6)) as o:
        prev = None
        for nc, npts in sizes:
            for h in (e, o):
                if prev is None:
                    h.set_problem(sub_problem(ba, Q, nc, npts))
                else:
                    h.append(**delta(Q, prev[0], prev[1], nc, npts))
            se, so = e.solve(), o.solve()
            te, to = e.trajectory(), o.trajectory()
            assert len(te) == len(to) and se.termination == so.termination
            for a, b in zip(te, to):
                assert a.step_is_successful == b.step_is_successful and abs(a.cost - b.cost) <= 1e-6 * abs(b.cost)
            for a, b in zip(e.get_state(), o.get_state()):
                assert np.linalg.norm(a - b) <= 1e-5 * np.linalg.norm(b)
            prev = (nc, npts)


@pytest.mark.gpu
def test_engine_append_moves_only_new_parameters(ba, engine_lib):
    """Persistent device state: ba_append uploads the new poses/points and the (re-merged) observation arrays, not the old parameters."""
    Q, sizes = staged(ba, "trafalgar", [200, 257])
    (nc0, np0), (nc1, np1) = sizes
    with ba.Engine(max_num_iterations=2) as e:
        e.set_problem(sub_problem(ba, Q, nc0, np0))
        e.solve()
        h0, _ = engine_lib.io_bytes()
        e.append(**delta(Q, nc0, np0, nc1, np1))
        h1, _ = engine_lib.io_bytes()
    with ba.Engine(max_num_iterations=2) as f:
        g0, _ = engine_lib.io_bytes()
        f.set_problem(sub_problem(ba, Q, nc1, np1))
        g1, _ = engine_lib.io_bytes()
    assert (g1 - g0) - (h1 - h0) == 8 * (6 * nc0 + 4 + 3 * np0)


@pytest.mark.gpu
def test_engine_warm_start_radius(ba):
    P = ba.synth("ladybug", seed=2)
    with ba.Engine(max_num_iterations=5, warm_start_radius=1) as e:
        e.set_problem(P)
        e.solve()
        r_end = e.trajectory()[-1].radius
        state = e.get_state()
        e.append()                       # nothing new: the next solve starts where the last one stopped, radius included
        e.solve()
        t_warm = e.trajectory()
    assert t_warm[0].radius == r_end and r_end != 1e4
    Q = ba.Problem(pose=state[0], intr=state[1], cam_group=P.cam_group, cam_fixed=P.cam_fixed, pt=state[2], obs_cam=P.obs_cam, obs_pt=P.obs_pt, obs_uv=P.obs_uv)
    with ba.Engine(max_num_iterations=5, initial_trust_region_radius=r_end) as f:
        f.set_problem(Q)
        f.solve()
        t_ref = f.trajectory()
    assert [r.cost for r in t_warm] == [r.cost for r in t_ref]
