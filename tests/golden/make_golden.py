"""Generate tests/golden/*.npz from the REFERENCE's own device source (oracle/_ref/*.so =
/root/reference/cuda_src/deviceCode.cu + cuda_include/** compiled unmodified as host C++).

Run in the build container (the reference is not present on the GPU box):
    python tests/golden/make_golden.py
The reference ships no tests or golden vectors; these are the pins.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _oracle  # noqa: E402
from optix_renderer_b200 import scenes  # noqa: E402

KEEP = ["position", "normal", "uv", "albedo", "alpha", "materialID", "ltc", "sto_direct", "sto_no_vis", "accum",
        "rgba8", "hit_prim", "hit_uv"]


def unit_vectors(o, seed=7):
    rng = np.random.default_rng(seed)
    g = {}
    # RNG streams (cuda_include/lcg_random.cuh)
    cases = [(10007, 0, 0, 1024, 1024), (10007, 3, 5, 512, 512), (10008, 1919, 1079, 1920, 1080), (10070, 77, 9, 64, 64)]
    g["rng_cases"] = np.array(cases, np.int64)
    st, raw, flt = [], [], []
    for c in cases:
        s, r, f = o.u_rng(*c, 8)
        st.append(s); raw.append(r); flt.append(f)
    g["rng_state"], g["rng_raw"], g["rng_flt"] = np.array(st, np.uint32), np.array(raw), np.array(flt)
    # clipPolygon: random unit triangles/quads straddling the horizon, every mask reachable
    V = rng.normal(size=(400, 5, 3)).astype(np.float32)
    V /= np.linalg.norm(V, axis=2, keepdims=True)
    vc_in = np.where(np.arange(400) % 2 == 0, 3, 4).astype(np.int32)
    outs, ns = [], []
    for i in range(400):
        v = V[i].copy()
        if vc_in[i] == 3:
            v[3] = v[0]; v[4] = v[0]
        else:
            v[4] = v[0]
        V[i] = v
        n, vo = o.u_clip(int(vc_in[i]), v)
        ns.append(n); outs.append(vo)
    g["clip_in"], g["clip_vc"], g["clip_n"], g["clip_out"] = V, vc_in, np.array(ns, np.int32), np.array(outs)
    # integrateEdge
    A = rng.normal(size=(200, 2, 3)).astype(np.float32)
    A /= np.linalg.norm(A, axis=2, keepdims=True)
    g["edge_in"] = A
    g["edge_out"] = np.array([o.u_integrate_edge(a[0], a[1]) for a in A], np.float32)
    # fetchLtcMat + matrixInverse
    at = np.stack([rng.uniform(0.01, 1, 100), rng.uniform(0, 1.55, 100)], 1).astype(np.float32)
    ms, amps, invs = [], [], []
    for a, t in at:
        m, amp = o.u_fetch_ltc(float(a), float(t))
        ms.append(m); amps.append(amp); invs.append(o.u_mat_inverse(m))
    g["fetch_in"], g["fetch_m"], g["fetch_amp"], g["minv"] = at, np.array(ms), np.array(amps, np.float32), np.array(invs)
    # orthonormalBasis
    N = rng.normal(size=(100, 3)).astype(np.float32)
    N /= np.linalg.norm(N, axis=1, keepdims=True)
    N[0] = (0, 0, -1); N[1] = (0, 0, 1)
    g["onb_in"] = N
    g["onb_out"] = np.array([o.u_onb(n)[0] for n in N])
    # sample_direction / evaluate / pdf
    WO = rng.normal(size=(100, 3)).astype(np.float32)
    WO /= np.linalg.norm(WO, axis=1, keepdims=True)
    U = rng.uniform(0, 1, (100, 2)).astype(np.float32)
    B = rng.uniform(0.1, 0.9, (100, 3)).astype(np.float32)
    AL = rng.uniform(0.01, 1, 100).astype(np.float32)
    wi, pdf, ev, pd = [], [], [], []
    for i in range(100):
        w, p = o.u_sample_direction(WO[i], float(U[i, 0]), float(U[i, 1]), B[i], 1.0, float(AL[i]))
        wi.append(w); pdf.append(p)
        ev.append(o.u_evaluate(w, WO[i], B[i], float(AL[i])))
        pd.append(o.u_pdf(w, WO[i], B[i], 1.0, float(AL[i])))
    g["bsdf_wo"], g["bsdf_u"], g["bsdf_base"], g["bsdf_alpha"] = WO, U, B, AL
    g["bsdf_wi"], g["bsdf_pdf"], g["bsdf_eval"], g["bsdf_pdf2"] = (np.array(wi), np.array(pdf, np.float32), np.array(ev),
                                                                  np.array(pd, np.float32))
    # samplePointOnTriangle, pdfA2W, make_rgba
    T = rng.normal(size=(50, 3, 3)).astype(np.float32)
    TU = rng.uniform(0, 1, (50, 2)).astype(np.float32)
    g["tri_in"], g["tri_u"] = T, TU
    g["tri_out"] = np.array([o.u_sample_tri(T[i, 0], T[i, 1], T[i, 2], float(TU[i, 0]), float(TU[i, 1])) for i in range(50)])
    P = np.abs(rng.normal(size=(50, 3))).astype(np.float32)
    P[0, 2] = 0.0; P[1, 2] = 1e-9
    g["a2w_in"] = P
    g["a2w_out"] = np.array([o.u_pdf_a2w(float(p[0]), float(p[1]), float(p[2])) for p in P], np.float32)
    Cc = rng.uniform(-0.5, 1.5, (50, 3)).astype(np.float32)
    Cc[0] = (np.nan, 0.5, 2.0)
    g["rgba_in"] = Cc
    g["rgba_out"] = np.array([o.u_make_rgba(c) for c in Cc], np.uint32)
    return g


def frames(scene_fn, name, W, H, depth=4):
    sc = scene_fn(W, H)
    out = {}
    for kind, mode in (("reference-literal", 0), ("reference-intended", 1)):
        o = _oracle.Oracle(kind)
        o.load_scene(sc)
        assert o.set_options(mode, 4, depth if mode else 1) == 0
        for rt, rname in ((_oracle.LTC_BASELINE, "ltc"), (_oracle.RATIO, "ratio"), (_oracle.PATH, "path")):
            f0 = o.render(W, H, rt, 0)
            f1 = o.render(W, H, rt, 1, accum=f0["accum"])
            for k in KEEP:
                out["%s_m%d_a0_%s" % (rname, mode, k)] = f0[k]
            out["%s_m%d_a1_accum" % (rname, mode)] = f1["accum"]
            out["%s_m%d_a1_rgba8" % (rname, mode)] = f1["rgba8"]
            out["%s_m%d_rays" % (rname, mode)] = np.array([f0["rays"], f1["rays"]], np.int64)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print("wrote", name, sum(v.nbytes for v in out.values()) // 1024, "KiB raw")


def ltc_gbuffer_case():
    """BASELINE config 1 shape at reduced size: saved G-buffer of the floor/blocker scene + the
    reference's LTC over it + the ratio arithmetic inputs."""
    W = H = 48
    sc = scenes.scene_floor_blocker(W, H)
    o = _oracle.Oracle("reference-intended")
    o.load_scene(sc)
    o.set_options(1, 4, 1)
    f = o.render(W, H, _oracle.RATIO, 0)
    ltc = o.ltc_gbuffer(f["position"], f["normal"], f["albedo"], f["alpha"], sc["camera"]["pos"])
    np.savez_compressed(os.path.join(HERE, "gbuffer48.npz"), position=f["position"], normal=f["normal"],
                        albedo=f["albedo"], alpha=f["alpha"], materialID=f["materialID"], ltc_frame=f["ltc"],
                        ltc_gbuffer=ltc, sto_direct=f["sto_direct"], sto_no_vis=f["sto_no_vis"],
                        cam=np.asarray(sc["camera"]["pos"], np.float32))
    print("wrote gbuffer48")


if __name__ == "__main__":
    assert _oracle.available("reference-literal"), "build oracle/_ref first (make -C oracle ref)"
    o = _oracle.Oracle("reference-literal")
    o.load_scene(scenes.scene_floor_blocker(8, 8))  # LUT for fetchLtcMat
    np.savez_compressed(os.path.join(HERE, "units.npz"), **unit_vectors(o))
    print("wrote units")
    frames(scenes.scene_floor_blocker, "frames_floor32", 32, 32)
    frames(scenes.scene_cornell, "frames_cornell24", 24, 24)
    frames(scenes.scene_textured, "frames_textured24", 24, 24)  # RGBA8 diffuse textures (hit_miss.cuh:39-43)
    ltc_gbuffer_case()
