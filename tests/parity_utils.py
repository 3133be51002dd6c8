"""Shared parity checks: CUDA engine vs the CPU oracle (tests only)."""
import json
import os
from typing import NamedTuple

import numpy as np
import torch

import rq_oracle as O

REL_TOL = 1e-3   # north_star: loss, quantized embeddings and gradients within 1e-3 relative
TIE_REL = 1e-6   # north_star: ids bit-exact unless the fp32 top-2 gap is under 1e-6 relative


def rel_err(a, b):
    a = torch.as_tensor(a).double().cpu()
    b = torch.as_tensor(b).double().cpu()
    denom = b.abs().max().clamp_min(1e-30)
    return float((a - b).abs().max() / denom)


def assert_close(a, b, what, tol=REL_TOL):
    e = rel_err(a, b)
    assert e <= tol, f"{what}: max rel err {e:.3e} > {tol}"


class IdCheck(NamedTuple):
    """Result of check_ids.  Unpacks as (n_mismatch_rows, n_outside_band_def1) for the older call sites."""
    n_mismatch_rows: int
    n_outside_band_def1: int


_RATES_LOG = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out", "parity_rates.jsonl")


def report_rates(what, rows, n_mismatch, n_out1, n_out2=0, **extra):
    """Mismatch rates of a parity run: printed (pytest -s / failure output) and appended to
    gpurun_out/parity_rates.jsonl when that directory exists (copied to profiles/ per round)."""
    rec = {"case": what, "rows": int(rows), "id_mismatch_rows": int(n_mismatch),
           "mismatch_rate": (n_mismatch / rows) if rows else 0.0,
           "outside_band_def1": int(n_out1), "outside_band_def2": int(n_out2)}
    rec.update(extra)
    print("[parity]", json.dumps(rec))
    d = os.path.dirname(_RATES_LOG)
    if os.path.isdir(d):
        with open(_RATES_LOG, "a") as f:
            f.write(json.dumps(rec) + "\n")
    return rec


def check_ids(ids_gpu, oracle_out, codebooks, what=""):
    """ids_gpu [L,B] vs oracle; a row may differ only from a level whose fp32 top-2 gap is inside
    the tie band (both definitions of SURVEY.md §7.3-3 are evaluated; definition 2 — relative to
    ||x||^2+||c||^2, the error scale of the expanded form — is the gate).  Returns
    IdCheck(n_mismatch_rows, n_outside_band_def1) and reports the rates (report_rates)."""
    ids_gpu = torch.as_tensor(ids_gpu).cpu()
    ids_ref = oracle_out.sem_ids
    Lv, B = ids_ref.shape
    diff = ids_gpu != ids_ref
    rows = torch.nonzero(diff.any(dim=0)).flatten()
    out1 = 0
    for r in rows.tolist():
        l = int(torch.nonzero(diff[:, r]).flatten()[0])
        x = oracle_out.residuals[l, r:r + 1]
        in1, in2, gap, d1 = O.tie_band(x, codebooks[l], TIE_REL)
        # the engine's choice must itself be within the band of the true minimum (with duplicated or
        # equidistant codewords there can be more than two such codes)
        d = ((x.double() - codebooks[l].double()) ** 2).sum(1)
        scale2 = float((x.double() ** 2).sum() + (codebooks[l][int(ids_gpu[l, r])].double() ** 2).sum())
        excess = float(d[int(ids_gpu[l, r])] - d.min())
        assert excess < TIE_REL * scale2, (f"{what}: row {r} level {l}: id {int(ids_gpu[l, r])} is {excess:.3e} "
                                           f"farther than the nearest code (band {TIE_REL * scale2:.3e})")
        assert bool(in2[0]), (f"{what}: row {r} level {l}: ids differ ({int(ids_gpu[l, r])} vs "
                              f"{int(ids_ref[l, r])}) outside the tie band: gap {float(gap[0]):.3e}, d1 {float(d1[0]):.3e}")
        out1 += int(not bool(in1[0]))
    # outside definition 2 would have failed the assert above, so that count is 0 whenever we get here
    report_rates(what, B, len(rows), out1, 0, levels=int(Lv))
    return IdCheck(len(rows), out1)


def matching_rows(ids_gpu, oracle_out):
    """Boolean [B]: rows whose ids agree at every level (float outputs are compared on these)."""
    return (torch.as_tensor(ids_gpu).cpu() == oracle_out.sem_ids).all(dim=0)


def synthetic_codebooks(latents: torch.Tensor, L: int, K: int, seed: int = 0):
    """Codebooks that behave like trained ones: level l = K residual rows (+ jitter) of the data,
    so assignments are spread and residual norms shrink level by level."""
    g = torch.Generator().manual_seed(seed)
    res = latents.clone()
    cbs = []
    for l in range(L):
        idx = torch.randperm(res.shape[0], generator=g)[:K]
        if idx.numel() < K:
            idx = torch.randint(0, res.shape[0], (K,), generator=g)
        cb = res[idx] * 0.9 + 0.02 * res.std() * torch.randn(K, res.shape[1], generator=g)
        cbs.append(cb.contiguous())
        out = O.quantize_level(res, cb, O.MODE_EVAL)
        res = res - out.emb_out
    return cbs
