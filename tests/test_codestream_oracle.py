"""Pins oracle/codestream.py (the checker of the GPU code-stream tests) against the reference itself, imported
unmodified from /root/reference (skipped where the reference is absent, e.g. on the GPU box), and checks the host
mirror audiocraft_b200/codes.py (layouts, lengths, argument checks) without a GPU."""
import importlib
import os
import sys
import types

import numpy as np
import pytest
import torch

from oracle import codestream as ocs
from tests.golden import gen

REF = "/root/reference"


def ref_patterns():
    """audiocraft.modules.codebooks_patterns without running audiocraft/modules/__init__.py (it needs xformers)."""
    path = os.path.join(REF, "audiocraft", "modules", "codebooks_patterns.py")
    if not os.path.isfile(path):
        pytest.skip("reference not available")
    from oracle import ref_shim
    ref_shim.load()  # registers the `audiocraft` package shell
    if "audiocraft.modules" not in sys.modules:
        pkg = types.ModuleType("audiocraft.modules")
        pkg.__path__ = [os.path.join(REF, "audiocraft", "modules")]
        sys.modules["audiocraft.modules"] = pkg
    return importlib.import_module("audiocraft.modules.codebooks_patterns")


CASES = [(4, 50, None, 0), (8, 300, None, 0), (4, 33, [0, 0, 0, 0], 0), (4, 20, [0, 1, 1, 3], 2), (3, 5, [0, 2, 7], 0),
         (2, 1, None, 0)]


@pytest.mark.parametrize("K,T,delays,empty_initial", CASES)
def test_delay_oracle_equals_reference(K, T, delays, empty_initial):
    cp = ref_patterns()
    prov = cp.DelayedPatternProvider(K, delays=delays, empty_initial=empty_initial)
    pat = prov.get_pattern(T)
    z = gen.randint(900 + K * T, (3, K, T), 2048)
    special = 2048
    offset = 1 + empty_initial
    assert [[(c.t, c.q) for c in step] for step in pat.layout] == ocs.delay_layout(prov.delays, offset, T)
    for keep in (False, True):
        v, i, m = pat.build_pattern_sequence(torch.from_numpy(z), special, keep_only_valid_steps=keep)
        ov, oi, om = ocs.delay_build(z, prov.delays, offset, special, keep)
        assert np.array_equal(v.numpy(), ov) and np.array_equal(i.numpy(), oi) and np.array_equal(m.numpy(), om)
        r, ri, rm = pat.revert_pattern_sequence(v, special, keep_only_valid_steps=keep)
        orr, ori, orm = ocs.delay_revert(ov, prov.delays, offset, special, T, keep)
        assert np.array_equal(r.numpy(), orr) and np.array_equal(ri.numpy(), ori) and np.array_equal(rm.numpy(), orm)
    # a partial sequence
    v, _, _ = pat.build_pattern_sequence(torch.from_numpy(z), special)
    part = v[:, :, : max(1, v.shape[2] // 2)].contiguous()
    r, ri, rm = pat.revert_pattern_sequence(part, special)
    orr, ori, orm = ocs.delay_revert(part.numpy(), prov.delays, offset, special, T, False)
    assert np.array_equal(r.numpy(), orr) and np.array_equal(ri.numpy(), ori) and np.array_equal(rm.numpy(), orm)


def test_parallel_pattern_is_the_zero_delay_case():
    cp = ref_patterns()
    pat = cp.ParallelPatternProvider(4).get_pattern(10)
    assert [[(c.t, c.q) for c in step] for step in pat.layout] == ocs.delay_layout([0] * 4, 1, 10)


@pytest.mark.parametrize("per_timestep", [False, True])
def test_stereo_oracle_equals_the_reference_rearrange(per_timestep):
    """models/encodec.py:476-479, 483-486 use exactly these einops patterns."""
    from einops import rearrange
    c0 = torch.from_numpy(gen.randint(31, (3, 4, 17), 2048))
    c1 = torch.from_numpy(gen.randint(32, (3, 4, 17), 2048))
    indices = torch.stack([c0, c1], dim=0)
    want = rearrange(indices, 'c b k t -> b k (t c)', c=2) if per_timestep else rearrange(indices, 'c b k t -> b (k c) t', c=2)
    got = ocs.stereo_interleave(c0.numpy(), c1.numpy(), per_timestep)
    assert np.array_equal(got, want.numpy())
    back = rearrange(want, 'b k (t c) -> c b k t', c=2) if per_timestep else rearrange(want, 'b (k c) t -> c b k t', c=2)
    l, r = ocs.stereo_split(got, per_timestep)
    assert np.array_equal(l, back[0].numpy()) and np.array_equal(r, back[1].numpy())


def test_pack_bits_oracle_round_trip():
    c = gen.randint(77, (1003,), 2048)
    p = ocs.pack_bits(c, 11)
    assert p.size == (1003 * 11 + 7) // 8
    bits = np.unpackbits(p, bitorder="little")[: 1003 * 11].reshape(1003, 11)
    assert np.array_equal((bits.astype(np.int64) << np.arange(11)).sum(1), c)


def test_host_mirror_layout_and_checks():
    from audiocraft_b200 import codes as cs
    prov = cs.DelayedPatternProvider(4, delays=[0, 1, 1, 3], empty_initial=2)
    pat = prov.get_pattern(20)
    assert [[tuple(c) for c in step] for step in pat.layout] == ocs.delay_layout([0, 1, 1, 3], 3, 20)
    assert len(pat.valid_layout) == 3 + 20 and pat.max_delay == 3 and pat.starts_with_special_token()
    assert cs.DelayedPatternProvider(3, empty_initial=-1).get_pattern(5).offset == 0
    assert cs.ParallelPatternProvider(4).delays == [0, 0, 0, 0]
    with pytest.raises(AssertionError):
        cs.DelayedPatternProvider(3, delays=[2, 1, 0])
    with pytest.raises(NotImplementedError):
        cs.DelayedPatternProvider(3, flatten_first=2)
    with pytest.raises(RuntimeError):        # CUDA tensors only: no CPU fallback
        pat.build_pattern_sequence(torch.zeros(1, 4, 20, dtype=torch.long), 2048)
    assert cs.bits_for(2048) == 11 and cs.bits_for(1024) == 10 and cs.bits_for(400) == 9
