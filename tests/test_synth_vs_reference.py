"""magritte_b200.synth (numpy construction of the hot-path inputs, used by bench.py on the GPU box where no Model
host exists yet) against the inputs the compiled reference derives for the same spec: bit for bit."""
import numpy as np
import pytest

import util
from magritte_b200 import flatten, modelbuild as mb, synth


@pytest.mark.parametrize("name", ["c1_all_constant", "velgrad_1d", "wind3d_spiral_2spec"])
def test_flat_from_spec_matches_reference(refcore_mod, name):
    core, _ = refcore_mod.load()
    spec = util.small_specs()[name]
    with util.quiet():
        m = mb.build(core, core.IoPython("mem", "synth_" + name), "synth_" + name, spec)
        mb.prepare(m)
    ref = flatten.flatten_model(m)
    got = synth.flat_from_spec(spec)
    for k, v in ref.items():
        if k == "params":
            assert got["params"] == v
            continue
        g = got[k]
        if isinstance(v, np.ndarray):
            assert g.shape == v.shape and g.dtype == v.dtype, k
            assert np.array_equal(g, v), f"{k}: max rel err {util.relerr(g, v)}"
        else:
            assert g == v, k


def test_golden_inputs_match_synth():
    """Without the reference: the committed golden inputs equal what synth builds today."""
    flat, _ = util.load_golden("wind3d_small")
    got = synth.flat_from_spec(util.small_specs()["wind3d_small"])
    for k in ("position", "neighbors", "emissivity", "opacity", "inverse_width", "nu", "population_tot", "antipod"):
        assert np.array_equal(got[k], flat[k]), k
