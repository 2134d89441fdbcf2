"""Shared parity checks: the same assertions run against the product's device headers compiled
for the host (tests/test_hostcheck.py, CPU) and against the CUDA kernels through the C-ABI
(tests/test_gpu_*.py).  `dut` = device under test with the u_* / render interface of
optix_renderer_b200.api.Renderer; golden = tests/golden/units.npz (from the reference source).

Tolerances (written here once):
  integer / index / RNG work ............ bit-exact
  single fp32 device functions .......... <= 4 ulp-ish (rtol 5e-7 + tiny atol): CUDA libdevice
                                          sqrt/div are IEEE, sinf/cosf/acosf differ from glibc
                                          by <= 2 ulp, and nvcc contracts a*b+c into fma
  LTC radiance per pixel ................ <= 1e-4 relative (BASELINE.json north_star)
  stochastic images ..................... RMSE <= 1e-3 (north_star), per-pixel error divided by
                                          max(1, |reference|) so bright pixels count relatively
"""
import os

import numpy as np

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
RT = 2e-6
AT = 2e-7


def units():
    return np.load(os.path.join(G, "units.npz"))


def close(a, b, rtol=RT, atol=AT):
    np.testing.assert_allclose(np.asarray(a, np.float64), np.asarray(b, np.float64), rtol=rtol, atol=atol)


def check_rng(dut, u):
    for c, s, raw, flt in zip(u["rng_cases"], u["rng_state"], u["rng_raw"], u["rng_flt"]):
        fr, px, py, W, _ = [int(x) for x in c]
        s2, r2, f2 = dut.u_rng(fr, px, py, W, 8)
        assert s2 == s and np.array_equal(r2, raw) and np.array_equal(f2, flt)  # bit-exact


def oracle_polygon_sum(ora, v, n):
    if n < 3:
        return 0.0
    s = np.float32(0)
    for i in range(n):
        s = np.float32(s + np.float32(ora.u_integrate_edge(v[i], v[(i + 1) % n])))
    return float(abs(s))


def check_clip(dut, ora, u):
    """clipPolygon + the closed edge sum (product API returns the clipped form factor)."""
    flips = 0
    for v, vc, n, out in zip(u["clip_in"], u["clip_vc"], u["clip_n"], u["clip_out"]):
        if vc != 3:
            continue  # a first clip always starts from a triangle light
        n2, ff = dut.u_clip(v[:3].reshape(9), quad=False)
        if min(abs(v[:3, 2])) < 1e-6:
            flips += n2 != n
            continue
        assert n2 == n
        close(ff, oracle_polygon_sum(ora, out, int(n)), rtol=2e-5, atol=2e-6)
    assert flips == 0


def check_scalar_units(dut, u):
    for a, r in zip(u["edge_in"], u["edge_out"]):
        close(dut.u_integrate_edge(a[0], a[1]), r, rtol=5e-6, atol=5e-7)
    for (al, th), m, amp, inv in zip(u["fetch_in"], u["fetch_m"], u["fetch_amp"], u["minv"]):
        m2, a2 = dut.u_fetch_ltc(float(al), float(th))
        close(m2, m, rtol=5e-6, atol=1e-6)
        close(a2, amp, rtol=5e-6)
        close(dut.u_mat_inverse(m), inv, rtol=1e-5, atol=1e-6)
    for n, r in zip(u["onb_in"], u["onb_out"]):
        close(dut.u_onb(n), r)
    for t, uu, r in zip(u["tri_in"], u["tri_u"], u["tri_out"]):
        close(dut.u_sample_tri(t.reshape(9), float(uu[0]), float(uu[1])), r, rtol=1e-6, atol=1e-6)
    for p, r in zip(u["a2w_in"], u["a2w_out"]):
        close(dut.u_pdf_a2w(float(p[0]), float(p[1]), float(p[2])), r)
    for c, r in zip(u["rgba_in"], u["rgba_out"]):
        assert dut.u_make_rgba(c) == r  # bit-exact


def check_bsdf(dut, u):
    for i in range(len(u["bsdf_alpha"])):
        wo, uu, b, al = u["bsdf_wo"][i], u["bsdf_u"][i], u["bsdf_base"][i], float(u["bsdf_alpha"][i])
        wi, p = dut.u_sample_direction(wo, float(uu[0]), float(uu[1]), b, al)
        close(wi, u["bsdf_wi"][i], rtol=1e-5, atol=1e-6)
        close(p, u["bsdf_pdf"][i], rtol=2e-5, atol=1e-7)
        close(dut.u_evaluate(u["bsdf_wi"][i], wo, b, al), u["bsdf_eval"][i], rtol=1e-5)
        close(dut.u_pdf(u["bsdf_wi"][i], wo, b, al), u["bsdf_pdf2"][i], rtol=2e-5, atol=1e-7)


def rel_err(got, ref, floor=1e-3):
    """per-pixel max relative error over rgb, |ref| floored (dark pixels are judged absolutely)."""
    got, ref = np.asarray(got, np.float64)[..., :3], np.asarray(ref, np.float64)[..., :3]
    return (np.abs(got - ref) / np.maximum(np.abs(ref), floor)).max(axis=-1)


def frame_report(got, ref, hit_got, hit_ref):
    """Parity summary of one frame against the oracle (north_star's reporting rules):
    hit-id mismatches are counted; radiance compared on pixels whose primary hit agrees."""
    same = hit_got == hit_ref
    out = {"pixels": int(same.size), "hit_mismatch": int((~same).sum())}
    for k in got:
        g, r = got[k], ref[k]
        nan_g, nan_r = np.isnan(g[..., :3]).any(-1), np.isnan(r[..., :3]).any(-1)
        m = same & ~nan_g & ~nan_r
        e = rel_err(g[m], r[m]) if m.any() else np.zeros(1)
        d = (np.asarray(g[m], np.float64)[..., :3] - np.asarray(r[m], np.float64)[..., :3])
        # RMSE with bright pixels judged relatively: a path that ends on a light after a low-pdf
        # bounce carries radiance in the hundreds, where 1e-3 ABSOLUTE would be 1e-6 relative
        d = d / np.maximum(1.0, np.abs(np.asarray(r[m], np.float64)[..., :3]))
        out[k] = {"max_rel": float(e.max()), "frac_le_1e-4": float((e <= 1e-4).mean()),
                  "rmse": float(np.sqrt((d * d).mean())) if d.size else 0.0,
                  "nan_mismatch": int((nan_g != nan_r)[same].sum()), "nan_ref": int(nan_r.sum())}
    return out


# ---- classification of the pixels that differ (north_star: "edge ties ... are counted and reported") -------------
def _prim_vertices(scene, prim):
    """(3, 3) float64 corner positions of global primitive `prim` (meshes in scene order, orb200.h)."""
    for m in scene["meshes"]:
        n = m["idx"].shape[0]
        if prim < n:
            return np.asarray(m["vtx"], np.float64)[m["idx"][prim]]
        prim -= n
    raise IndexError(prim)


def _ray_segment_angle(c, d, p, q):
    """Smallest angle (radians, small-angle form) under which the segment [p, q] is seen from the ray c + s d."""
    e, w = q - p, c - p
    b, cc, dd, f = d @ e, e @ e, d @ w, e @ w
    den = cc - b * b
    u = 0.0 if den <= 1e-300 else min(1.0, max(0.0, (f - b * dd) / den))
    s = b * u - dd
    return float(np.linalg.norm(w + s * d - u * e) / max(s, 1e-12))


def classify_hit_mismatches(got, ref, cam_pos, scene=None, edge_tol=1e-4, angle_tol=5e-6):
    """Split the pixels whose primary hit id differs into
      tie    both sides hit, at the same point in space (|dp| <= 16 ulp of the hit distance): two triangles share
             the edge / the point the ray passes through and the closest-t comparison went the other way;
      edge   the ray grazes the boundary of the NEARER of the two triangles -- the one the other side missed: seen
             from the camera it passes within `angle_tol` radians (1/120 of a 1080p pixel, 1/60 of a 4K pixel; measured
             worst case 3.3e-6 on the 4K / 10 M-triangle frame, silhouette triangles seen edge-on) of one of its edges, so a last-ulp difference of the camera ray (fma contraction) moves it on
             or off the triangle and what lies behind is hit instead.  Without `scene` (no vertices): the hit lies
             within `edge_tol` (barycentric) of an edge on either side;
      other  anything else: a real difference.  Tests fail on these.
    got / ref need hit_prim (H, W), hit_uv (H, W, 2) and position (H, W, 4)."""
    gp, rp = got["hit_prim"], ref["hit_prim"]
    bad = gp != rp
    out = {"mismatch": int(bad.sum()), "tie": 0, "edge": 0, "other": 0}
    if not bad.any():
        return out
    cam = np.asarray(cam_pos, np.float64)
    pg = np.asarray(got["position"], np.float64)[bad][:, :3]
    pr = np.asarray(ref["position"], np.float64)[bad][:, :3]
    g, r = gp[bad], rp[bad]
    ug, ur = got["hit_uv"][bad].astype(np.float64), ref["hit_uv"][bad].astype(np.float64)
    eps = float(np.finfo(np.float32).eps)
    worst = 0.0
    for k in range(len(g)):
        tg = np.linalg.norm(pg[k] - cam) if g[k] >= 0 else np.inf
        tr = np.linalg.norm(pr[k] - cam) if r[k] >= 0 else np.inf
        if g[k] >= 0 and r[k] >= 0 and np.abs(pg[k] - pr[k]).max() <= 16 * eps * max(1.0, tr):
            out["tie"] += 1
            continue
        near_prim, near_p = (g[k], pg[k]) if tg <= tr else (r[k], pr[k])
        if scene is not None:
            d = (near_p - cam) / np.linalg.norm(near_p - cam)
            v = _prim_vertices(scene, int(near_prim))
            ang = min(_ray_segment_angle(cam, d, v[i], v[(i + 1) % 3]) for i in range(3))
            worst = max(worst, ang)
            is_edge = ang <= angle_tol
        else:
            def near(uv, prim):
                return prim >= 0 and min(uv[0], uv[1], 1.0 - uv[0] - uv[1]) <= edge_tol
            is_edge = near(ug[k], g[k]) or near(ur[k], r[k])
        out["edge" if is_edge else "other"] += 1
        if not is_edge:
            out.setdefault("others", []).append({"got": int(g[k]), "ref": int(r[k]), "t_got": float(tg), "t_ref": float(tr),
                                                 "edge_angle_rad": float(ang) if scene is not None else None})
    if scene is not None:
        out["max_edge_angle_rad"] = worst
    return out


def classify_ltc_outliers(err, margin, same_hit, tol=1e-4, near=1e-3):
    """Pixels whose LTC radiance is outside `tol` (relative), split by the distance of their evaluation to its
    nearest sign decision (oracle ora_ltc_margin: wo_local.z, back-face test, above-horizon bits of the clip):
      flip   margin <= near -- the pixel takes, or sits beside, another case of clipPolygon's table;
      other  margin >  near -- a real difference.  Tests fail on these."""
    bad = same_hit & (err > tol)
    flip = bad & (margin <= near)
    return {"pixels": int(same_hit.sum()), "outside_tol": int(bad.sum()), "flip": int(flip.sum()),
            "other": int((bad & ~flip).sum()),
            "margin_quartiles_of_outliers": [float(x) for x in np.quantile(margin[bad], [0.25, 0.5, 0.75, 1.0])] if bad.any() else []}
