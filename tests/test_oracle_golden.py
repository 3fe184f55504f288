"""CPU: pin the oracle (oracle/ecolab_oracle.cc) to the reference's own golden vectors and tests.

Sources (paths in the reference checkout):
  test/00/predator-prey.sh:18-44 input, :56-67 golden   -- generate + ROUND + mt19937 consumption order
  test/00/condense.sh:14-21                              -- condense packs [1,0,1,0,1] -> [1,1,1]
  test/00/spatial_ecolab.sh:43-77                        -- generate changes every cell; migrate conserves species totals
  test/testarrays.cc:284-327                             -- float->int truncation, pack, enumerate, gen_index, sparse mat-vec
"""
import json
import os

import numpy as np
import pytest

from helpers import densities, load_oracle, synth

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")

PREDATOR_PREY = [(96, 79), (99, 102), (108, 81), (101, 95), (96, 83), (103, 101), (101, 79), (102, 102), (102, 83), (97, 98)]


def predator_prey(o):
    o.seed(10)
    o.set("species", [1, 2]); o.set_density(0, [100, 100]); o.set("create", [0, 0]); o.set("repro_rate", [.1, -.1])
    o.set("interaction.diag", [-.0001, -1e-5]); o.set("interaction.val", [-0.001, 0.001])
    o.set("interaction.row", [0, 1]); o.set("interaction.col", [1, 0])
    o.set("mutation", [.01, .01]); o.set("sp_sep", .1); o.set("repro_min", -.1); o.set("repro_max", .1)
    o.set("odiag_min", -1e-3); o.set("odiag_max", 1e-3); o.set("mut_max", .01)


def test_predator_prey_known_answer(oracle_mod):
    O = oracle_mod
    o = O.Oracle(rng=O.RNG_MT19937)
    predator_prey(o)
    got = []
    for _ in range(10):
        for _ in range(100):
            o.generate()
        o.condense()
        got.append(tuple(int(v) for v in o.get_density(0)))
    assert got == PREDATOR_PREY


def test_predator_prey_generate_100_equals_100_generates(oracle_mod):
    O = oracle_mod
    a, b = O.Oracle(rng=O.RNG_MT19937), O.Oracle(rng=O.RNG_MT19937)
    predator_prey(a); predator_prey(b)
    a.generate(100)
    for _ in range(100):
        b.generate()
    assert list(a.get_density(0)) == list(b.get_density(0)) == list(PREDATOR_PREY[0])


def test_condense_sh(oracle_mod):
    O = oracle_mod
    o = O.Oracle(rng=O.RNG_MT19937)
    o.set_density(0, [1, 0, 1, 0, 1])
    o.make_consistent()
    o.set("repro_rate", [1] * 5); o.set("interaction.diag", [1] * 5)
    assert o.condense() == 2
    assert list(o.get_density(0)) == [1, 1, 1]
    assert list(o.get("species")) == [0, 2, 4]


def test_condense_remaps_interaction(oracle_mod):
    # ModelData::condense :139-159: keep entry iff both ends alive; remap through enumerate(mask); order preserved
    O = oracle_mod
    o = O.Oracle(rng=O.RNG_PHILOX)
    o.set("species", range(5)); o.set("repro_rate", [0.] * 5); o.set("interaction.diag", [-1., -2., -3., -4., -5.])
    o.set("interaction.row", [4, 0, 1, 2, 4, 3]); o.set("interaction.col", [0, 4, 2, 4, 2, 0])
    o.set("interaction.val", [10., 11., 12., 13., 14., 15.])
    o.set_density(0, [5, 0, 7, 0, 9]); o.make_consistent()
    assert o.condense() == 2
    assert list(o.get("interaction.row")) == [2, 0, 1, 2]
    assert list(o.get("interaction.col")) == [0, 2, 2, 1]
    assert list(o.get("interaction.val")) == [10., 11., 13., 14.]
    assert list(o.get("interaction.diag")) == [-1., -3., -5.]


def spatial_sh_model(O, rng_mode):
    rs = np.random.default_rng(3)
    nsp = 30
    o = O.Oracle(rng=rng_mode, seed=1)
    o.set("species", range(nsp)); o.set("repro_min", -.1); o.set("repro_max", .1); o.set("odiag_min", -1e-5); o.set("odiag_max", 1e-5)
    o.set("repro_rate", rs.uniform(-.1, .1, nsp)); o.set("interaction.diag", np.full(nsp, -1e-3))
    O.srand(1); O.reset_grand()
    o.random_interaction(3, 0)
    nnz = len(o.get("interaction.row"))
    o.set("interaction.val", rs.uniform(-1e-5, 1e-5, nnz))
    o.set("mut_max", 1e-4); o.set("sp_sep", 0.1); o.set("mutation", np.full(nsp, 1e-4)); o.set("migration", np.full(nsp, 1e-5))
    o.set_grid(2, 2)
    for c in range(4):
        o.set_density(c, [100] * nsp)
    return o, nsp


@pytest.mark.parametrize("mode", ["mt19937", "philox"])
def test_spatial_ecolab_sh(oracle_mod, mode):
    O = oracle_mod
    o, nsp = spatial_sh_model(O, O.RNG_MT19937 if mode == "mt19937" else O.RNG_PHILOX)
    o.generate()
    d = o.get_density_all()
    assert all((d[c] != 100).any() for c in range(4))
    n0 = d.sum(axis=0)
    o.migrate()
    assert (o.get_density_all().sum(axis=0) == n0).all()
    # migration 1e-5 * 1 step on ~100 individuals truncates to zero flux: give it a real gradient
    o.set("migration", np.full(nsp, 0.1))
    o.set_density(0, [1000] * nsp)
    n0 = o.get_density_all().sum(axis=0)
    o.generate()
    n0 = o.get_density_all().sum(axis=0)
    moved = o.migrate()
    d1 = o.get_density_all()
    assert moved > 0 and (d1.sum(axis=0) == n0).all() and (d1 >= 0).all()


def test_random_interaction_shape(oracle_mod):
    # sparse_mat::init_rand, include/sparse_mat.h:88-111: conn distinct columns per row, self links dropped, val zeroed
    O = oracle_mod
    o, nsp = spatial_sh_model(O, O.RNG_PHILOX)
    row, col = o.get("interaction.row"), o.get("interaction.col")
    assert (row != col).all()
    assert (np.diff(row.astype(np.int64)) >= 0).all()
    for i in range(nsp):
        c = col[row == i]
        assert len(c) in (2, 3) and len(set(c.tolist())) == len(c)


def test_testarrays_primitives(oracle_mod):
    O = oracle_mod
    # float -> int assignment truncates toward zero (test/testarrays.cc:284-290): migrate's (int)(cm*(nb-n))
    o = O.Oracle(rng=O.RNG_PHILOX)
    o.set("species", [0]); o.set("repro_rate", [0.]); o.set("interaction.diag", [0.]); o.set("migration", [0.3]); o.set("mutation", [0.])
    o.set_grid(2, 1)
    o.set_density(0, [10]); o.set_density(1, [3])
    o.set("tstep", 1)
    assert o.get("maxNbrs") == 1
    moved = o.migrate()          # cm = min(0.3, 1/1); trunc(0.3*-7) = -2, trunc(0.3*7) = 2
    assert list(o.get_density_all().ravel()) == [8, 5] and moved == 2
    # sparse mat-vec = diag*x then scatter row += val*x[col] (test/testarrays.cc:318-327) through generate's pre-rounding x
    o = O.Oracle(rng=O.RNG_PHILOX)
    o.set("species", [0, 1, 2]); o.set("repro_rate", [0., 0., 0.]); o.set("interaction.diag", [1., 2., 3.])
    o.set("interaction.row", [0, 2, 0]); o.set("interaction.col", [1, 0, 2]); o.set("interaction.val", [.5, .25, .125])
    o.set_density(0, [2, 4, 8]); o.set("record_pre", 1)
    o.generate()
    x = o.pre_rounding()
    ir = [1 * 2 + .5 * 4 + .125 * 8, 2 * 4, 3 * 8 + .25 * 2]
    assert list(x) == [2 + 2 * ir[0], 4 + 4 * ir[1], 8 + 8 * ir[2]]
    # gen_index (test/testarrays.cc:311-316) through mutate: parents repeated `count` times, density appended
    o = O.Oracle(rng=O.RNG_PHILOX)
    o.set("species", [0, 1]); o.set("repro_rate", [1., 1.]); o.set("interaction.diag", [-1., -1.]); o.set("create", [0, 0])
    o.set("mutation", [1., 1.]); o.set("migration", [0., 0.]); o.set("sp_sep", 1.0); o.set("mut_max", 1.0)
    o.set_density(0, [2, 0]); o.set("tstep", 1)
    O.srand(5); O.reset_grand()
    o.mutate()                   # mut_scale = 1 -> ROUND(2.0) = 2 mutants of species 0, none of species 1
    assert list(o.get_density(0)) == [0, 0, 1, 1]
    assert list(o.get("species")) == [0, 1, 2, 3]


def test_round_edges(oracle_mod):
    O = oracle_mod
    R = O.round_with_draw
    assert R(-1.5, 0) == 0 and R(0.0, 0) == 0
    assert R(2147483646.0, 0) == 2147483646 and R(1e300, 0) == 2147483646
    assert R(5.0, 0) == 5 and R(5.0, 0xFFFFFFFF) == 5
    assert R(5.5, 0) == 6 and R(5.5, 0xFFFFFFFF) == 5
    assert R(5.5, 2147483647) == 6 and R(5.5, 2147483648) == 5   # 0.5*4294967295 = 2147483647.5
    assert R(1e-12, 0) == 1 and R(1e-12, 1) == 0


def test_philox_known_answer(oracle_mod):
    # Random123 kat_vectors, philox4x32-10
    O = oracle_mod
    assert O.philox_raw([0, 0, 0, 0], [0, 0]) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert O.philox_raw([0xffffffff] * 4, [0xffffffff] * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert O.philox_raw([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0]) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]
    # the contract: ctr = {(x>>2) | stage<<30, y, slot, t_lo}, key = {seed, t_hi}, draw = block[x & 3]
    blk = O.philox_raw([(9 >> 2) | (1 << 30), 7, 5, 3], [11, 2])
    assert O.philox(11, 1, 9, 7, 5, 3 + (2 << 32)) == blk[9 & 3]
    assert [O.philox(11, 1, x, 7, 5, 3 + (2 << 32)) for x in (8, 9, 10, 11)] == blk
    base = O.philox(1, 0, 2, 6, 4, 3)
    assert len({base, O.philox(2, 0, 2, 6, 4, 3), O.philox(1, 1, 2, 6, 4, 3), O.philox(1, 0, 3, 6, 4, 3), O.philox(1, 0, 6, 6, 4, 3),
                O.philox(1, 0, 2, 7, 4, 3), O.philox(1, 0, 2, 6, 5, 3), O.philox(1, 0, 2, 6, 4, 4),
                O.philox(1, 0, 2, 6, 4, 3 + (1 << 32))}) == 9


def test_committed_golden_fixtures(oracle_mod):
    """tests/golden/*.npz were produced by tests/golden/make_golden.py (oracle, Philox mode); the oracle must still
    reproduce them -- they are also what the GPU tests compare against on the box."""
    O = oracle_mod
    from golden.make_golden import CASES, run_case
    assert CASES
    for name in CASES:
        want = np.load(os.path.join(GOLDEN, name + ".npz"))
        got = run_case(O, name)
        for k in want.files:
            assert np.array_equal(want[k], got[k]), (name, k)
