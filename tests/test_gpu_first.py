"""First GPU parity tests: sampler words and tallies, bit-exact against the oracle, through the C ABI."""
import numpy as np
import pytest

from helpers import O, random_graph
from canopy_b200 import synth
from canopy_b200.engine import Plan

pytestmark = pytest.mark.gpu


def test_sampler_words_bit_exact():
    g = synth.generate(200, 100, 5, 8, seed=11)
    with Plan(g) as plan:
        v = plan.view()
        got = plan.sample_words(seed=0xDEADBEEFCAFE, first_word=64 * 5, n_words=128)
    want = O.sample_words(g, 0xDEADBEEFCAFE, v["event_orig"], 64 * 5, 128)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("seed", range(8))
def test_random_graph_tallies_bit_exact(seed):
    g = random_graph(seed, n_basic=14, n_gates=40, n_roots=3)
    with Plan(g, tally_all_gates=True) as plan:
        got, _ = plan.run(10_000, seed=seed + 99, sample_begin=123)
    assert np.array_equal(got, O.mc_scalar(g, seed + 99, 123, 10_000, True))
