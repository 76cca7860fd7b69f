"""Parity checker: CUDA path (through the C-ABI) vs the CPU oracle on the same inputs.

Rule (BASELINE.json north_star): accepted hypothesis index sets and support counts bit-exact; any disagreement must be a
near-threshold case (oracle comparison margin < 1e-9) and is counted and reported; 3-D locations and tangents within
1e-6 relative.  Intermediate FP64 values may differ by the libm difference (glibc vs CUDA libdevice, <= 2 ulp per call).
"""
import numpy as np

MARGIN = 1e-9      # a disagreement is tolerated only if the oracle recorded a comparison margin below this
REL_GEOM = 1e-6    # 3-D points / tangents
ABS_ANGLE = 1e-10  # degrees; epipolar angles and wedge bounds (pure libm noise is ~1e-13)
REL_ROWS = 1e-9    # corrected / clustered 2-D locations


def _relerr(a, b):
    d = np.abs(a - b)
    s = np.maximum(np.abs(b), 1e-300)
    return float(np.max(d / np.maximum(s, 1.0))) if a.size else 0.0


def compare(orc, gpu, check_intermediates=False, check_finalize=True):
    """orc: oracle_py.Oracle after run(keep=True); gpu: api.EdgeSketch after run().  Returns a report dict; raises
    AssertionError on a violation of the parity rule."""
    rep = {"threshold_disagreements_form": 0, "threshold_disagreements_support": 0, "argmax_ties": 0}
    # ---- pass-level tables
    a_o, a_g = orc.h2_angles(), gpu.h2_angles()
    assert np.array_equal(np.isnan(a_o), np.isnan(a_g))
    rep["h2_angle_maxabs"] = float(np.nanmax(np.abs(a_o - a_g))) if a_o.size else 0.0
    assert rep["h2_angle_maxabs"] <= ABS_ANGLE
    w_o, f_o = orc.h1_wedges()
    w_g, f_g = gpu.h1_wedges()
    assert np.array_equal(f_o, f_g), "view-1 wedge flags differ"
    ok = f_o == 0
    rep["h1_wedge_maxabs"] = float(np.max(np.abs(w_o[ok] - w_g[ok]))) if ok.any() else 0.0
    assert rep["h1_wedge_maxabs"] <= ABS_ANGLE
    # ---- per view-1 edge tables
    t_o, t_g = orc.h1_table(), gpu.h1_table()
    fm = t_o["form_margin"]
    r_o, r_g = orc.rows(), gpu.rows()
    s_o, s_g = orc.support_matrix(), gpu.support_matrix()
    inter = orc.intermediates() if orc.keep else None
    E1 = len(t_o["m"])
    bad_h1 = set()
    for h in range(E1):
        same = (t_o["m"][h] == t_g["m"][h]) and (t_o["n"][h] == t_g["n"][h])
        if same and t_o["n"][h] > 0:
            so, sg = t_o["off"][h], t_g["off"][h]
            n = t_o["n"][h]
            for k in ("h2_orig", "local_idx", "raw_at_local", "label", "h2_rep"):
                if not np.array_equal(r_o[k][so:so + n], r_g[k][sg:sg + n]):
                    same = False
        if not same:
            assert fm[h] < MARGIN, f"view-1 edge {h}: hypothesis table differs and the oracle margin is {fm[h]:.3e}"
            bad_h1.add(h)
    rep["threshold_disagreements_form"] = len(bad_h1)
    rep["bad_h1"] = sorted(int(h) for h in bad_h1)      # view-1 edges excluded downstream (oracle margin < 1e-9), for exact set comparisons
    # ---- rows: values and support matrix
    maxrel = 0.0
    n_sup_diff = 0
    n_checked = 0
    for h in range(E1):
        n = t_o["n"][h]
        if n == 0 or h in bad_h1:
            continue
        so, sg = t_o["off"][h], t_g["off"][h]
        for k in ("xc", "yc", "theta_orig", "xbar", "ybar", "thbar"):
            maxrel = max(maxrel, _relerr(r_g[k][sg:sg + n], r_o[k][so:so + n]))
        d = s_o[so:so + n] != s_g[sg:sg + n]
        n_checked += d.size
        if d.any():
            assert inter is not None, f"support differs for view-1 edge {h} (run the oracle with keep=True to attribute it)"
            mg = inter["margin"][so:so + n][d]
            assert (mg < MARGIN).all(), f"support differs for view-1 edge {h} with oracle margins {mg}"
            n_sup_diff += int(d.sum())
    rep["rows_maxrel"] = maxrel
    assert maxrel <= REL_ROWS, f"row values differ by {maxrel:.3e}"
    rep["threshold_disagreements_support"] = n_sup_diff
    rep["validations_compared"] = n_checked
    # ---- counts
    c_o, c_g = orc.counts(), gpu.counts()
    rep["counts_oracle"], rep["counts_gpu"] = c_o, c_g
    exact = (not bad_h1) and n_sup_diff == 0
    if exact:
        for k in ("n_h1_processed", "n_candidates", "n_hypotheses", "n_validations", "n_valid_rows", "n_supported", "n_parallel", "n_inliers", "n_unique_hypotheses",
                  "n_degenerate_wedges"):
            assert c_o[k] == c_g[k], f"count {k}: oracle {c_o[k]} gpu {c_g[k]}"
    # ---- vote / selection, compared per view-1 edge on (h1, row within the edge); edges with a near-threshold
    # disagreement upstream are excluded (and were counted above)
    skip = set(bad_h1)
    if n_sup_diff:
        for h in range(E1):
            n = t_o["n"][h]
            if n and h not in bad_h1 and (s_o[t_o["off"][h]:t_o["off"][h] + n] != s_g[t_g["off"][h]:t_g["off"][h] + n]).any():
                skip.add(h)
    def _keyed(h1, row, off, extra):
        keep = np.array([h not in skip for h in h1.tolist()], bool) if skip else np.ones(len(h1), bool)
        loc = row.astype(np.int64) - off[h1]
        return h1[keep], loc[keep], [e[keep] for e in extra]

    v_o, v_g = orc.valid_rows(), gpu.valid_rows()
    ho, lo_, eo = _keyed(v_o["h1"], v_o["row"], t_o["off"], [v_o["count"]])
    hg, lg_, eg = _keyed(v_g["h1"], v_g["row"], t_g["off"], [v_g["count"]])
    assert np.array_equal(ho, hg) and np.array_equal(lo_, lg_) and np.array_equal(eo[0], eg[0]), "valid rows (count >= N) differ"
    # accepted rows in the reference's paired_edge order: std::sort by count + ratio flag + per-edge capacity (host step of the ABI)
    p_g = gpu.pairs()
    acc_idx = gpu.reference_order()
    p_o = orc.pairs()
    ao_h, ao_l, ao_e = _keyed(p_o["h1"], p_o["row"], t_o["off"], [p_o["h2rep"], p_o["count"], p_o["supports"], np.arange(len(p_o["h1"]))])
    ag_h, ag_l, ag_e = _keyed(p_g["h1"][acc_idx], p_g["row"][acc_idx], t_g["off"],
                              [p_g["h2rep"][acc_idx], p_g["count"][acc_idx], p_g["supports"][acc_idx], acc_idx])
    assert np.array_equal(ao_h, ag_h) and np.array_equal(ao_l, ag_l), "accepted rows (reference order) differ"
    assert np.array_equal(ao_e[0], ag_e[0]), "representative view-2 edge of accepted pairs differs"
    assert np.array_equal(ao_e[1], ag_e[1]) and np.array_equal(ao_e[2], ag_e[2]), "support counts / indices of accepted pairs differ"
    rep["n_pairs_compared"] = int(len(ao_h))
    rep["n_pairs_oracle"] = int(len(p_o["h1"]))
    if check_finalize and len(ao_h):
        f_o, f_g = orc.finalize(), gpu.finalize()
        io, ig = ao_e[3], ag_e[3]
        assert np.array_equal(f_o["nan_flag"][io], f_g["nan_flag"][ig])
        for k in ("gamma_cam", "gamma_world", "tangent_world"):
            a, b = f_g[k][ig], f_o[k][io]
            scale = np.maximum(np.linalg.norm(b, axis=1, keepdims=True), 1e-12)
            err = float(np.max(np.abs(a - b) / scale))
            rep[k + "_maxrel"] = err
            assert err <= REL_GEOM, f"{k} differs by {err:.3e} relative"
        rep["h2_corrected_maxabs"] = float(np.nanmax(np.abs(f_g["h2_corrected"][ig] - f_o["h2_corrected"][io])))
        assert rep["h2_corrected_maxabs"] <= 1e-8
    if check_intermediates and inter is not None and not bad_h1:
        gi = gpu.intermediates()
        proc = np.nonzero(t_o["n"] > 0)[0]
        w31g = gi["wedge31"][proc]
        both = ~np.isnan(inter["wedge31"]) & ~np.isnan(w31g)
        assert np.array_equal(np.isnan(inter["wedge31"]), np.isnan(w31g))
        rep["wedge31_maxabs"] = float(np.max(np.abs(inter["wedge31"][both] - w31g[both]))) if both.any() else 0.0
        assert np.array_equal(np.isnan(inter["wedge32"]), np.isnan(gi["wedge32"]))
        both = ~np.isnan(inter["wedge32"])
        rep["wedge32_maxabs"] = float(np.max(np.abs(inter["wedge32"][both] - gi["wedge32"][both]))) if both.any() else 0.0
        both = ~np.isnan(inter["tgt"]) & ~np.isnan(gi["tgt"])
        rep["tgt_maxabs"] = float(np.max(np.abs(inter["tgt"][both] - gi["tgt"][both]))) if both.any() else 0.0
        assert rep["wedge31_maxabs"] <= ABS_ANGLE and rep["wedge32_maxabs"] <= ABS_ANGLE and rep["tgt_maxabs"] <= 1e-9
    return rep


def assert_claims_equal(orc, masks_o, masks_g, bad_h1, view2):
    """Claimed-edge masks (one uint8 array per view): exact, except that view-2 bits of the hypothesis rows of `bad_h1` (view-1 edges
    whose cluster representative is a sub-1e-9 tie) may differ.  Returns the number of differing bits (all attributed)."""
    t, r = orc.h1_table(), orc.rows()
    unsure = set()
    for h in bad_h1:
        n, s = int(t["n"][h]), int(t["off"][h])
        unsure.update(r["h2_orig"][s:s + n].tolist())
        unsure.update(r["raw_at_local"][s:s + n].tolist())
    n_diff = 0
    for v, (mo, mg) in enumerate(zip(masks_o, masks_g)):
        d = np.nonzero(np.asarray(mo, bool) != np.asarray(mg, bool))[0]
        if v == view2:
            assert set(d.tolist()) <= unsure, f"view {v}: claim bits differ outside the near-tie rows: {sorted(set(d.tolist()) - unsure)[:10]}"
        else:
            assert len(d) == 0, f"view {v}: {len(d)} claim bits differ"
        n_diff += len(d)
    return n_diff
