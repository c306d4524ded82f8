"""Basis-conversion workload for timing / ncu captures.
Usage: python tools/prof_basis.py [logN] [sizeQ] [sizeP] [batch] [reps]
Times ApproxSwitchCRTBasis Q->P, ApproxModUp, ApproxModDown, DropLastElementAndScale on device-resident operands."""
import importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("upmem--openfhe_b200")
logn = int(sys.argv[1]) if len(sys.argv) > 1 else 16
sq = int(sys.argv[2]) if len(sys.argv) > 2 else 30
sp = int(sys.argv[3]) if len(sys.argv) > 3 else 10
batch = int(sys.argv[4]) if len(sys.argv) > 4 else 16
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 5
n = 1 << logn


def timed(be, f):
    for _ in range(2):
        f()
    be.sync()
    be.timer_start(0)
    for _ in range(reps):
        f()
    return be.timer_stop(0) / reps


with pkg.Backend(0, "arena_mb=1024") as be:
    q, psi = pkg.gen_params(60, n, sq + sp)
    ring = pkg.Ring(be, n, q, psi)
    rq, rp = pkg.RingView(ring, 0, sq), pkg.RingView(ring, sq, sp)
    q2p, p2q, plan = pkg.Basis(rq, rp), pkg.Basis(rp, rq), pkg.Rescale(ring)
    wq, wp = batch * sq * n, batch * sp * n
    dq, dp, dqp, dr = be.alloc(wq * 8), be.alloc(wp * 8), be.alloc((wq + wp) * 8), be.alloc((wq + wp) * 8)
    rq.fill_uniform(dq, batch, 1)
    ring.fill_uniform(dqp, batch, 2)
    out = {"N": n, "sizeQ": sq, "sizeP": sp, "batch": batch}
    ms = timed(be, lambda: q2p.approx_switch(dp, dq, batch))
    out["approx_switch_q_to_p"] = {"ms": ms, "mac128_per_s": batch * n * sq * sp / (ms * 1e-3),
                                   "coeff_in_per_s": wq / (ms * 1e-3), "hbm_gbs": 8 * (wq + wp) / (ms * 1e-3) / 1e9}
    ms = timed(be, lambda: p2q.approx_switch(dq, dp, batch))
    out["approx_switch_p_to_q"] = {"ms": ms, "mac128_per_s": batch * n * sq * sp / (ms * 1e-3),
                                   "hbm_gbs": 8 * (wq + wp) / (ms * 1e-3) / 1e9}
    rq.fill_uniform(dq, batch, 1)
    ms = timed(be, lambda: q2p.approx_mod_up(dr, dq, batch, 0))
    out["approx_mod_up_eval"] = {"ms": ms, "polys_per_s": batch / (ms * 1e-3)}
    ms = timed(be, lambda: p2q.approx_mod_down(dq, dqp, batch, 0))
    out["approx_mod_down"] = {"ms": ms, "polys_per_s": batch / (ms * 1e-3)}
    ms = timed(be, lambda: plan.drop_last_element_and_scale(dr, dqp, batch, 0))
    out["drop_last_element_and_scale"] = {"ms": ms, "polys_per_s": batch / (ms * 1e-3),
                                          "towers": sq + sp}
    # HPSPOVERQ tails (tower_scale_kernel): 8 B in + 8 B out per coefficient per tower
    T1 = sq + sp - 1
    rq1 = pkg.RingView(ring, 0, T1)
    ts = pkg.TowerScale(rq1, [int(v) // 3 for v in q[:T1]])
    ms = timed(be, lambda: ts.scale_and_round_p_over_q(dr, dqp, batch))
    out["scale_and_round_p_over_q"] = {"ms": ms, "hbm_gbs": batch * n * (2 * T1 + 1) * 8 / (ms * 1e-3) / 1e9, "towers": T1}
    ms = timed(be, lambda: ts.expand_crt_basis_ql_hat(dr, dq, sq, batch))
    out["expand_crt_basis_ql_hat"] = {"ms": ms, "hbm_gbs": batch * n * (sq + T1) * 8 / (ms * 1e-3) / 1e9, "towers": T1}
    ts.close()
    rq1.close()
    print(json.dumps(out))
    for o in (q2p, p2q, plan, rq, rp):
        o.close()
    ring.close()
