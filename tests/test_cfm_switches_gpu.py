"""Every A/B switch of the estimator selects another kernel path for the SAME math: the fused cluster kernels against their
separate-GEMM forms, the tcgen05 attention against the mma.sync one, the N-split LayerNorm GEMMs against split-K, the
feed-forward's exchange through L2 against distributed shared memory.  The switches are read once per process, so each variant
is a 2-step solve at L = 300 in its own process (tools/cfm_once.py); all variants must agree with the default build to the
rounding of a different summation order (bf16 activations flip by an ulp here and there: measured 3.4e-3 relative L2 after two
steps for every variant; bound 1e-2), and the L2 / DSMEM
forms of the feed-forward exchange bit for bit.

The padded-batch kernels are only reached by shape (>= 592 GEMM tiles, > 296 attention CTAs): a second test runs a 2-step solve of
16 utterances at L = 812 (tools/cfm_batch_time.py) with the weight-stationary GEMM off, with the one-query-tile attention kernel
in place of the two-tile one, and with the two-group attention kernel forced; the first two must agree bit for bit (same
arithmetic, other schedule), the third to rounding."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
VARIANTS = {"outffn0": {"CBX_CFM_OUTFFN": "0"}, "ffn0": {"CBX_CFM_FFN": "0"}, "red_dsmem": {"CBX_FFN_RED_L2": "0"},
            "nsplit1": {"CBX_GEMM_NSPLIT": "1"}, "attn_mma": {"CBX_ATTN_TC": "0"}, "nograph": {"CBX_CFM_GRAPH": "0"}}


def _solve(tmp_path, name, env):
    out = str(tmp_path / f"{name}.pt")
    e = dict(os.environ, CFM_SAVE=out, CFM_REPS="1", **env)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "cfm_once.py"), "300", "2"], env=e, capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return torch.load(out)


def test_every_estimator_switch_computes_the_same_solve(tmp_path):
    base = _solve(tmp_path, "default", {})
    assert torch.isfinite(base).all()
    for name, env in VARIANTS.items():
        got = _solve(tmp_path, name, env)
        rel = float((got - base).norm() / base.norm())
        print(f"{name:10s} {env}: rel-L2 vs default {rel:.2e}")
        assert rel < 1e-2, name
        if name in ("red_dsmem", "nograph"):
            assert torch.equal(got, base), name


BATCH_VARIANTS = {"gemm_ws0": {"CBX_GEMM_WS": "0"}, "attn_qt0": {"CBX_ATTN_QT": "0"}, "attn_groups2": {"CBX_ATTN_GROUPS": "2"}}


def _solve_batch(tmp_path, name, env):
    out = str(tmp_path / f"b_{name}.pt")
    e = dict(os.environ, CFM_SAVE=out, CFM_REPS="1", CFM_STEPS="2", **env)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "cfm_batch_time.py"), "16", "812"], env=e, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return torch.load(out)


def test_padded_batch_kernels_agree_with_their_fallbacks(tmp_path):
    base = _solve_batch(tmp_path, "default", {})
    assert torch.isfinite(base).all()
    for name, env in BATCH_VARIANTS.items():
        got = _solve_batch(tmp_path, name, env)
        rel = float((got - base).norm() / base.norm())
        print(f"{name:12s} {env}: rel-L2 vs default {rel:.2e}")
        assert rel < 1e-2, name
        if name in ("gemm_ws0", "attn_qt0"):
            assert torch.equal(got, base), name
