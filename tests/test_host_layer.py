"""CPU tests of the host-side C++ layer (prophet_b200/host): Optimizer, checkpoint I/O, Setup, readers."""
import os

import numpy as np
import pytest

from golden_util import GOLDEN_DIR, TRAIN_CASES, TrainCase
from prophet_b200 import hostapi, synth


def _ensure_built():
    if not os.path.exists(hostapi.LIB_PATH):
        import __graft_entry__ as g
        g.build()


@pytest.mark.parametrize("name", TRAIN_CASES)
def test_rprop_reproduces_reference_steps_exactly(name):
    """Fed the reference's own gradients, Optimizer::update_network + rprop must return the reference's next
    parameter vector bit for bit (the optimizer is sign based: anything less diverges)."""
    _ensure_built()
    c = TrainCase(name)
    ntrain = int(c.d["Ntrain"][0])
    opt = hostapi.Optimizer(c.d["x_before.0"], "rprop", ntrain)
    for it in range(c.n_iter):
        x1 = opt.step(c.d["SSE.%d" % it][0], c.d["gradient.%d" % it])
        assert np.array_equal(x1, c.d["x_after.%d" % it]), (name, it)


@pytest.mark.parametrize("name,algorithm,kw", [
    ("opt_bfgs", "bfgs", dict(line_min_epsilon=1e-3)),
    ("opt_sd_line", "steepest_descent_with_line_opt", dict(line_min_epsilon=1e-3)),
    ("opt_default_switch", "default", dict(line_min_epsilon=1e-3, debug=0.5)),
])
def test_line_search_algorithms_reproduce_reference_calls_exactly(name, algorithm, kw):
    """L-BFGS + Brent line minimisation (Optimizer.cpp:607-869), steepest descent with line optimisation (:457-507)
    and the "default" schedule Rprop -> steepest descent -> L-BFGS (:305-322, :436-447): fed the SSE and gradient of
    every Optimizer::update_network call of a reference run, the parameter vector handed back must be the
    reference's bit for bit -- including the calls inside a line search, which do not advance the iteration
    counter, and (default schedule) the reference's own collapse to NaN once y.s = 0."""
    _ensure_built()
    d = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    n = sum(1 for k in d.files if k.startswith("gradient."))
    opt = hostapi.Optimizer(d["x_before.0"], algorithm, int(d["Ntrain"][0]), **kw)
    for it in range(n):
        x1 = opt.step(d["SSE.%d" % it][0], d["gradient.%d" % it])
        assert np.array_equal(x1, d["x_after.%d" % it], equal_nan=True), (name, it)
    if name == "opt_default_switch":
        # the schedule really left Rprop: plain Rprop on the same inputs departs from the reference at call 133
        rp = hostapi.Optimizer(d["x_before.0"], "rprop", int(d["Ntrain"][0]))
        same = [np.array_equal(rp.step(d["SSE.%d" % it][0], d["gradient.%d" % it]), d["x_after.%d" % it]) for it in range(140)]
        assert all(same[:133]) and not same[133]


def test_perturbation_scheme_follows_the_reference(capfd):
    """lambda_max != 0 (Optimizer.cpp:915-938, 1008-1043): reaching the threshold switches a slow sine perturbation on
    instead of stopping; the scheme decides when training ends and draws random numbers.  On the quadratic objective of
    the golden run (ref_driver optq) the parameters every 500 calls, the call at which the optimizer stops, the final
    parameters and the state of libc's rand() stream must all equal the reference's."""
    import ctypes
    _ensure_built()
    d = np.load(os.path.join(GOLDEN_DIR, "opt_perturbation.npz"))
    P = int(d["spec_P"]); seed = int(d["spec_seed"])
    libc = ctypes.CDLL("libc.so.6")
    libc.srand(seed)
    for _ in range(2 * P):
        libc.rand()                       # the reference's node constructor drew P normal deviates for the initial weights
    a = np.array([0.5 + 0.37 * ((i * 7) % 5) for i in range(P)])
    b = np.array([0.25 * ((i * 3) % 7) - 0.6 for i in range(P)])
    opt = hostapi.Optimizer(d["x0"], "rprop", 1)
    opt.configure(d["x0"], lambda_max=float(d["spec_lambda_max"]), ncycles=int(d["spec_ncycles"]),
                  threshold=float(d["spec_threshold"]), do_print=True)
    x = d["x0"].copy()
    stop = int(d["converged_at"][0])
    for call in range(1, stop + 1):
        f = 0.0                            # accumulated in the reference's order
        for i in range(P):
            f += a[i] * (x[i] - b[i]) * (x[i] - b[i])
        x = opt.step(f, 2 * a * (x - b))
        if call % 500 == 0:
            assert np.array_equal(x, d["x.%d" % call]), call
        assert opt.converged() == (call == stop), call
    assert np.array_equal(x, d["x_final"])
    assert libc.rand() == int(d["next_rand"][0])
    out = capfd.readouterr().out
    for msg in ("converged; turning on perturbation", "transition detected", "continuing optimization", "No transition found"):
        assert msg in out


def test_approximate_ks_statistic_matches_reference():
    """The appendix of `-validate` (Potential.cpp:1107-1117 -> Analysis::KS, Analysis.cpp:149-202)."""
    _ensure_built()
    d = np.load(os.path.join(GOLDEN_DIR, "ks.npz"))
    for k in range(4):
        D, p = hostapi.ks(d["a.%d" % k], d["b.%d" % k])
        assert D == d["D.%d" % k][0] and p == d["p.%d" % k][0], (k, D, p)
    assert hostapi.ks(d["a.3"], d["a.3"]) == (0.0, 1.0)


def test_default_algorithm_starts_as_rprop():
    _ensure_built()
    c = TrainCase("si_train")
    opt = hostapi.Optimizer(c.d["x_before.0"], "default", int(c.d["Ntrain"][0]))
    for it in range(c.n_iter):
        assert np.array_equal(opt.step(c.d["SSE.%d" % it][0], c.d["gradient.%d" % it]), c.d["x_after.%d" % it])


def test_steepest_descent_first_steps():
    """Optimizer.cpp:400-449: first step 5e-5 along the normalised gradient, then 1.4x / 0.7x step adaptation."""
    _ensure_built()
    x0 = np.array([1.0, -2.0, 0.5])
    opt = hostapi.Optimizer(x0, "steepest_descent", 1)
    g0 = np.array([3.0, 0.0, -4.0])
    x1 = opt.step(10.0, g0)
    assert np.allclose(x1, x0 - 5e-5 * g0 / 5.0, rtol=0, atol=1e-18)
    g1 = np.array([0.0, 2.0, 0.0])
    x2 = opt.step(9.0, g1)           # error went down -> alpha *= 1.4
    assert np.allclose(x2, x1 - 5e-5 * 1.4 * g1 / 2.0, rtol=0, atol=1e-18)
    x3 = opt.step(9.5, g1)           # error went up -> alpha *= 0.7
    assert np.allclose(x3, x2 - 5e-5 * 1.4 * 0.7 * g1 / 2.0, rtol=0, atol=1e-18)


@pytest.mark.parametrize("sym", ["O", "Ti"])
def test_checkpoint_roundtrip_is_byte_identical(sym, tmp_path):
    """A potential file written by the reference's Potential::save, read and written back by the host layer."""
    _ensure_built()
    src = os.path.join(GOLDEN_DIR, "ref_checkpoint_%s.txt" % sym)
    out = str(tmp_path / "copy")
    hostapi.checkpoint_roundtrip(src, out)
    assert open(out, "rb").read() == open(src, "rb").read()


def test_checkpoint_python_writer_agrees_with_cpp_reader(tmp_path):
    """The harness-side writer (prophet_b200/ckpt.py) and the C++ reader/writer agree on the format."""
    _ensure_built()
    from prophet_b200 import ckpt, synth
    cfg = synth.config_c2(1)
    net = synth.random_network(cfg, 14, 3)
    p = str(tmp_path / "pot_Si")
    ckpt.write_checkpoint(p, "Si", ["Si"], {"G1": [], "G2": cfg.G2, "G3": [], "G4": cfg.G4}, np.linspace(0, 1, cfg.NG),
                          np.linspace(1, 2, cfg.NG), -0.25, 1.5, net["sizes"], net["params"])
    hostapi.checkpoint_roundtrip(p, p + ".rt")
    a = ckpt.read_checkpoint(p)
    b = ckpt.read_checkpoint(p + ".rt")
    assert np.array_equal(a["params"], b["params"]) and a["sizes"] == b["sizes"] and a["node_transfer"] == b["node_transfer"]
    assert np.allclose(a["means"], b["means"], rtol=1e-11) and a["rows"] == b["rows"]


def test_setup_parses_reference_grammar_and_shards_round_robin(tmp_path):
    """Keys are case-insensitive, '=' optional, '#' comments, include files, data blocks, directory lines with
    train|val|test tags (Setup.cpp:135-552); systems are dealt round robin over ranks (Setup.cpp:582)."""
    _ensure_built()
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        for k in range(5):
            os.makedirs("data/s%d" % k)
        with open("systems", "w") as f:
            f.write("data code=vasp   # block\n./data/s0\n./data/s1 val\n./data/s2   train\n./data/s3 test\n./data/s4\n")
        with open("input_file", "w") as f:
            f.write("# comment\nNetwork_Type = nn\nhidden=  20 10\nINPUT= structure\noutput energy\ntransfer = tanh\n"
                    "G2 5.5 0 0.5 1.0\nG4 4.5 0 0.01 2 -1\nRcut=4\nalgorithm=rprop\nNiterations = 30\nNprint 5\n"
                    "checkpoint_out = pot\nNcheckpoint = 7\nseed = 3\ninclude systems\n")
        rep = hostapi.setup_parse("input_file")
        assert "hidden 20 10" in rep and "transfer tanh" in rep and "algorithm rprop" in rep
        assert "Niterations 30" in rep and "Nprint 5" in rep and "Ncheckpoint 7" in rep
        assert "rows 0 1 0 1" in rep and "checkpoint_out pot" in rep
        # Rcut is raised to the largest row cutoff by the first directory line (Setup.cpp:517-519)
        assert "Rcut 5.5" in rep
        assert "systems_total 5 local 5" in rep
        assert "energy=./data/s1" in rep and "structure=./data/s1" in rep and "train=val" in rep
        r0 = hostapi.setup_parse("input_file", 0, 2)
        r1 = hostapi.setup_parse("input_file", 1, 2)
        assert "local 3" in r0 and "local 2" in r1
        assert all(("structure=./data/s%d" % k) in r0 for k in (0, 2, 4))
        assert all(("structure=./data/s%d" % k) in r1 for k in (1, 3))
    finally:
        os.chdir(cwd)


def test_auto_generated_rows_match_reference():
    """Structure::init_G's default rows (Structure.cpp:818-851): the tutorial (Nradial=5 Nangular=4 Rcut=4) and the
    BaTiO3 fixture (Nradial=3 Nangular=4 Rcut=5) goldens hold what the reference generated."""
    _ensure_built()
    for name, rcut, nr, na in (("tutorial_c1", 4.0, 5, 4), ("batio3_doc", 5.0, 3, 4)):
        c = TrainCase(name)
        g2, g4 = hostapi.auto_rows(rcut, nr, na)
        assert np.array_equal(g2, c.G2) and np.array_equal(g4, c.G4), name


def test_vasprun_reader(tmp_path):
    """A vasprun.xml with the elements the reference's VASP reader looks for (VASP.cpp:78-176), written from the
    tutorial golden's own cell / positions / energy."""
    _ensure_built()
    c = TrainCase("tutorial_c1")
    d = tmp_path / "run1"
    d.mkdir()
    cell, frac, Z, target = c.cells[0], c.frac[0], c.Z[0], c.targets[0]
    synth.write_vasprun(str(d), cell, frac, Z, target)
    cell2, pos2, Z2, t2 = hostapi.read_vasprun(str(d))
    assert np.array_equal(cell2, cell) and np.array_equal(pos2, frac) and np.array_equal(Z2, Z) and t2 == target
