"""CPU-side checks: the C-ABI library loads and exports every symbol include/cbx.h declares, the
static plan covers every weight unit exactly once, and the packer lays units out as the plan says."""
import os
import re

import pytest
import torch

from chatterbox_b200 import _lib, pack
from chatterbox_b200 import synthetic as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "cbx.h")).read()
    declared = set(re.findall(r"\b(cbx_[a-z0-9_]+)\s*\(", hdr))
    l = _lib.lib()
    for name in declared:
        assert hasattr(l, name), name
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)
    assert l.cbx_version() >= 1


@pytest.mark.parametrize("G", [132, 148, 160])
def test_plan_is_a_partition(G):
    plan = pack.plan_lists(G)
    for ph, U in enumerate(pack.PHASE_UNITS):
        units = sorted(u for c in range(G) for u in plan[c][ph])
        assert units == list(range(U)), (G, ph)
        assert max(len(plan[c][ph]) for c in range(G)) <= 4
    # all DOWN units of a CTA share one k-slice
    for c in range(G):
        assert len({u % 4 for u in plan[c][pack.PH_DOWN]}) <= 1
    per_layer = [sum(len(plan[c][ph]) for ph in range(4)) for c in range(G)]
    assert max(per_layer) - min(per_layer) <= 2


def test_pack_layout_follows_plan():
    G, nl = 148, 1
    sd = S.make_state_dict("t3", n_layers=nl)
    pk = pack.pack_t3(sd, G, "cpu", max_pos=64, n_layers=nl)
    plan = pack.plan_lists(G)
    assert pk.wstream.shape[0] == G and pk.wstream.dtype == torch.bfloat16

    def unit_rows(c, slot):   # undo the A-fragment order -> [16 rows, 1024 k]
        f = pk.wstream[c, slot].float().reshape(pack.kwarps(), 64 // pack.kwarps(), 8, 4, 2, 2, 2)   # [w][s][g][tig][kh][rh][e]
        return f.permute(5, 2, 0, 1, 4, 3, 6).reshape(16, 1024)               # row = rh*8+g, k = w*16KS+s*16+kh*8+tig*2+e

    q = sd["tfmr.layers.0.self_attn.q_proj.weight"]
    gate, up = sd["tfmr.layers.0.mlp.gate_proj.weight"], sd["tfmr.layers.0.mlp.up_proj.weight"]
    down, head = sd["tfmr.layers.0.mlp.down_proj.weight"], sd["speech_head.weight"]
    bf = lambda t: t.to(torch.bfloat16).float()
    for c in (0, 57, 147):
        slot = 0
        u = plan[c][0][0]
        if u < 64:
            assert torch.equal(unit_rows(c, slot), bf(q[16 * u:16 * u + 16]))
        slot = len(plan[c][0]) + len(plan[c][1])
        u = plan[c][2][0]
        assert torch.equal(unit_rows(c, slot), bf(torch.cat([gate[8 * u:8 * u + 8], up[8 * u:8 * u + 8]])))
        slot += len(plan[c][2])
        if plan[c][3]:
            u = plan[c][3][-1]
            r, s = divmod(u, 4)
            assert torch.equal(unit_rows(c, slot + len(plan[c][3]) - 1), bf(down[16 * r:16 * r + 16, 1024 * s:1024 * s + 1024]))
        slot += len(plan[c][3])
        u = plan[c][4][0]
        assert torch.equal(unit_rows(c, slot), bf(head[16 * u:16 * u + 16]))
    # zero padding of the head beyond row 8194
    c = [c for c in range(G) if 512 in plan[c][4]][0]
    slot = nl * sum(len(plan[c][ph]) for ph in range(4)) + plan[c][4].index(512)
    rows = unit_rows(c, slot)
    assert torch.equal(rows[:2], bf(head[8192:8194])) and float(rows[2:].abs().max()) == 0.0


def test_rope_tables_match_oracle():
    from oracle import t3_oracle as T3O
    cos, sin = pack.rope_tables(300, "cpu")
    c2, s2 = T3O.rope_cos_sin(torch.arange(300))
    assert torch.equal(cos, c2[:, :32]) and torch.equal(sin, s2[:, :32])
