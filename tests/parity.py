"""Parity metrics between a GPU render (RenderResult) and an oracle render (dict) — BASELINE.json's bars:
hit ids agree on >= 99.99 % of primary rays with disagreements only at t ties within 1e-5 relative;
8-bit pixels within 1 LSB on >= 99.9 % of pixels."""
import numpy as np


def compare(got, ref, tie_rel=1e-5):
    rep = {}
    g_rgba = got.rgba if hasattr(got, "rgba") else got["rgba"]
    r_rgba = ref["rgba"]
    d = np.abs(g_rgba[..., :3].astype(np.int32) - r_rgba[..., :3].astype(np.int32)).max(axis=-1)
    rep["px_exact"] = float((d == 0).mean())
    rep["px_within_1"] = float((d <= 1).mean())
    rep["px_max_diff"] = int(d.max())
    rep["alpha_ok"] = bool((g_rgba[..., 3] == 255).all())
    g_ids = got.ids if hasattr(got, "ids") else got.get("ids")
    r_ids = ref.get("ids")
    if g_ids is not None and r_ids is not None:
        same = (g_ids["inst"] == r_ids["inst"]) & (g_ids["prim"] == r_ids["prim"]) & (g_ids["kind"] == r_ids["kind"])
        rep["id_agree"] = float(same.mean())
        dis = ~same
        rep["id_disagree"] = int(dis.sum())
        tg, tr = g_ids["t"][dis], r_ids["t"][dis]
        with np.errstate(invalid="ignore"):
            rel = np.abs(tg - tr) / np.maximum(np.abs(tr), 1e-300)
        tie = np.isfinite(tg) & np.isfinite(tr) & (rel <= tie_rel)
        rep["id_bad"] = int((~tie).sum())     # disagreements that are NOT t ties
        both = same & (r_ids["inst"] >= 0)
        rep["t_bit_exact"] = float((g_ids["t"][both] == r_ids["t"][both]).mean()) if both.any() else 1.0
    gs = got.stats if hasattr(got, "stats") else got.get("stats")
    if gs is not None and "counters" in ref:
        c = ref["counters"]
        rep["rays_gpu"] = (gs["raysPrimary"], gs["raysShadow"], gs["raysSecondary"])
        rep["rays_ref"] = (c["raysPrimary"], c["raysShadow"], c["raysSecondary"])
    return rep
