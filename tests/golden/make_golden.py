"""Regenerates tests/golden/*.json from the CPU oracle (oracle/canopy_oracle.c).

The reference (openpra-org/canopy) has no implementation, tests or golden vectors for this path, so these
fixtures pin the behaviour of the three sampling specs (1 "direct", 2 "sliced", 4 "classed sparse") as restated by the oracle -- they are NOT outputs of the reference.
What IS external: the three Philox4x32-10 known-answer vectors (Random123 kat_vectors, also listed in
SURVEY.md Appendix B), which are written here literally and checked against the oracle before anything
else is generated.

Run from the repository root:  python tests/golden/make_golden.py
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from helpers import O, random_graph  # noqa: E402
from canopy_b200 import synth  # noqa: E402
from canopy_b200.engine import load_mef  # noqa: E402

PHILOX_KAT = [
    dict(ctr=[0, 0, 0, 0], key=[0, 0], out=[0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]),
    dict(ctr=[0xffffffff] * 4, key=[0xffffffff] * 2, out=[0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]),
    dict(ctr=[0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], key=[0xa4093822, 0x299f31d0],
         out=[0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]),
]


def graph_dict(g):
    return dict(p=[float(x) for x in g.p], op=[int(x) for x in g.op], kmin=[int(x) for x in g.kmin],
                kmax=[int(x) for x in g.kmax], child_off=[int(x) for x in g.child_off],
                child=[int(x) for x in g.child], roots=[int(x) for x in g.roots])


def case(name, g, seed, begin, n, tally_all, bits=32, words_of=(0, 1, 2), spec=2):
    ev = [e for e in words_of if e < g.n_basic]
    return dict(name=f"{name}_spec{spec}", graph=graph_dict(g), seed=seed, begin=begin, n=n, tally_all=tally_all, bits=bits,
                spec=spec, tallies=[int(x) for x in O.mc_scalar(g, seed, begin, n, tally_all, bits, spec)],
                word_events=ev, first_word=begin // 64,
                words=[[int(w) for w in row] for row in O.sample_words(g, seed, ev, begin // 64, 2, bits, spec)])


def main():
    for k in PHILOX_KAT:
        assert O.philox4x32_10(k["ctr"], k["key"]) == k["out"], "oracle disagrees with the published Philox KAT"
    with open(os.path.join(HERE, "philox_kat.json"), "w") as f:
        json.dump(dict(source="Random123 kat_vectors, philox4x32-10; also SURVEY.md Appendix B", vectors=PHILOX_KAT), f, indent=1)
    cases = []
    c1 = load_mef(os.path.join(ROOT, "examples", "c1_small_ft.xml"))
    tiny = synth.generate(n_basic=120, n_gates=60, n_layers=5, max_fanin=16, seed=0xC0FFEE)
    for spec in (1, 2, 4):
        cases.append(case("c1_small_ft", c1, seed=0, begin=0, n=1 << 14, tally_all=True, spec=spec))
        cases.append(case("c1_small_ft_offset_seed", c1, seed=(0xDEADBEEF << 32) | 0x1234, begin=(1 << 40) + 77, n=5000,
                          tally_all=False, spec=spec))
        if spec == 4:   # graphs whose events cover the sparse classes k = 7 .. 13 and the dense boundary p = 2^-5 (the random
                        # graphs below are all dense); ranges straddle 8192-sample group boundaries; 32- and 64-bit thresholds
            for s in range(4):
                cases.append(case(f"random_rare_{s}", random_graph(3000 + s, n_basic=14, n_gates=30, n_roots=3, p_lo=1e-3, p_hi=0.06),
                                  seed=s * 104729 + 3, begin=128 * s + 5 * s, n=8192 + 123 * s, tally_all=True, spec=spec))
            for s in range(4):
                cases.append(case(f"random_classes_{s}", random_graph(4000 + s, n_basic=16, n_gates=30, n_roots=3, p_lo=5e-5, p_hi=0.02,
                                                                     extreme_p=s % 2 == 0),
                                  seed=s * 15485863 + 11, begin=8192 * (3 + s) - 700 * (s + 1), n=2 * 8192 + 77 * s, tally_all=True,
                                  bits=64 if s >= 2 else 32, spec=spec, words_of=(0, 1, 2, 3, 4, 5)))
        for s in range(6 if spec == 2 else 3):
            cases.append(case(f"random_{s}", random_graph(1000 + s, n_basic=12, n_gates=30, n_roots=3), seed=s * 7919 + 1,
                              begin=64 * s + s, n=4096 + 321 * s, tally_all=True, spec=spec))
        cases.append(case("random_thr64", random_graph(2000, n_basic=10, n_gates=20), seed=5, begin=3, n=3000, tally_all=True,
                          bits=64, spec=spec))
        cases.append(case("synth_tiny", tiny, seed=42, begin=4096 * 3 - 10, n=4096 + 20, tally_all=True, spec=spec))
    with open(os.path.join(HERE, "golden_tallies.json"), "w") as f:
        json.dump(dict(spec_versions=[1, 2, 4], generator="tests/golden/make_golden.py (oracle_mc_scalar / oracle_sample_words)",
                       cases=cases), f)
    print(f"wrote {len(cases)} cases")


if __name__ == "__main__":
    main()
