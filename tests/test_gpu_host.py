"""GPU tests of the drop-in surfaces: the PROPhet_b200 command line against the reference's own CLI on the same
input_file / PROPhet.xml / checkpoint files, and pair_style nn through the LAMMPS stand-in against the MD goldens."""
import os
import re
import subprocess

import numpy as np
import pytest

from golden_util import MDCase, rel_err
from prophet_b200 import ckpt, hostapi, synth

pytestmark = pytest.mark.gpu
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
REF_CLI = os.path.join(ROOT, "oracle", "_ref", "PROPhet")
IT_LINE = re.compile(r"^(\d+)\s+([-+0-9.eE]+)\s+([-+0-9.eE]+)(?:\s+([-+0-9.eE]+))?\s*$")


def _iteration_table(stdout):
    rows = []
    for line in stdout.splitlines():
        m = IT_LINE.match(line)
        if m:
            rows.append([float(x) for x in m.groups() if x is not None])
    return rows


def _write_case(cwd, cfg, n_struct, tags, hidden, seed, niter, extra=()):
    cfg.hidden = hidden
    st = [synth.make_structure(cfg, 40 + s) for s in range(n_struct)]
    st = [(c, x, z, t, tag) for (c, x, z, t, _), tag in zip(st, tags)]
    synth.write_prophet_xml(os.path.join(cwd, "PROPhet.xml"), st)
    syms = [synth.Z_TO_SYMBOL[z] for z in cfg.elements]
    for z in cfg.elements:
        net = synth.random_network(cfg, z, seed)
        ckpt.write_checkpoint(os.path.join(cwd, "start_" + synth.Z_TO_SYMBOL[z]), synth.Z_TO_SYMBOL[z], syms,
                              {"G1": cfg.G1, "G2": cfg.G2, "G3": cfg.G3, "G4": cfg.G4},
                              0.3 * np.abs(synth.normal(z, cfg.NG, 3)), 0.5 + np.abs(synth.normal(z, cfg.NG, 4)), 0.0, 1.0,
                              net["sizes"], net["params"], header_precision=17)
    synth.write_input_file(os.path.join(cwd, "input_file"), cfg, niterations=niter, checkpoint_in="start",
                           checkpoint_out="trained", extra=extra)


def _run(exe, args, cwd):
    r = subprocess.run([exe] + args, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:]
    return r.stdout


@pytest.mark.skipif(not os.path.exists(REF_CLI), reason="oracle/_ref/PROPhet not built")
@pytest.mark.parametrize("cfg_name,hidden", [("C2", [8, 6]), ("C3", [7, 5])])
def test_cli_train_matches_reference_cli(cfg_name, hidden, tmp_path):
    """Same input_file, PROPhet.xml and checkpoint_in for both programs: the printed training table and the saved
    potentials must agree (single- and two-element systems, with a validation structure -> early-stop table)."""
    assert os.path.exists(hostapi.CLI_PATH), "PROPhet_b200 not built"
    cfg = synth.CONFIGS[cfg_name](5)
    ours = tmp_path / "ours"; ref = tmp_path / "ref"
    for d in (ours, ref):
        d.mkdir()
        _write_case(str(d), cfg, 5, [0, 0, 1, 0, 2], hidden, seed=9, niter=6)
    out_ref = _run(REF_CLI, ["-in", "input_file", "-train"], str(ref))
    out_ours = _run(hostapi.CLI_PATH, ["-in", "input_file", "-train"], str(ours))
    t_ref, t_ours = _iteration_table(out_ref), _iteration_table(out_ours)
    assert len(t_ref) == 6 and len(t_ours) == 6, (out_ref[-1500:], out_ours[-1500:])
    assert np.allclose(np.array(t_ours), np.array(t_ref), rtol=2e-6, atol=0), (t_ours, t_ref)
    for z in cfg.elements:
        sym = synth.Z_TO_SYMBOL[z]
        a = ckpt.read_checkpoint(str(ours / ("trained_" + sym)))
        b = ckpt.read_checkpoint(str(ref / ("trained_" + sym)))
        assert a["sizes"] == b["sizes"] and a["rows"] == b["rows"] and a["node_transfer"] == b["node_transfer"]
        assert rel_err(a["params"], b["params"]) < 1e-9
        assert np.allclose(a["means"], b["means"], rtol=1e-11) and np.allclose(a["scales"], b["scales"], rtol=1e-11)
    # -validate and -run on the trained potentials: the prediction tables agree
    for d in (ours, ref):
        txt = open(str(d / "input_file")).read().replace("checkpoint_in = start", "checkpoint_in = trained")
        open(str(d / "input_file"), "w").write(txt)
    v_ref = _run(REF_CLI, ["-in", "input_file", "-validate"], str(ref))
    v_ours = _run(hostapi.CLI_PATH, ["-in", "input_file", "-validate"], str(ours))

    def table(txt):
        rows = []
        for line in txt.splitlines():
            tok = line.split()
            if len(tok) == 5 and tok[0].isdigit() and tok[4] in ("train", "val", "test", "other"):
                rows.append((int(tok[0]), float(tok[1]), float(tok[2]), int(tok[3]), tok[4]))
        return rows

    tr, to = table(v_ref), table(v_ours)
    assert len(tr) == 5 and len(to) == 5
    for a, b in zip(to, tr):
        assert a[0] == b[0] and a[3] == b[3] and a[4] == b[4]
        assert abs(a[1] - b[1]) <= 1e-8 * abs(b[1]) and a[2] == b[2]
    rms = lambda txt: float(re.search(r"RMS Error =\s+([-+0-9.eE]+)", txt).group(1))
    assert abs(rms(v_ours) - rms(v_ref)) <= 1e-5 * abs(rms(v_ref))
    # the approximate Kolmogorov-Smirnov appendix (Potential.cpp:1107-1117)
    ks = lambda txt: (float(re.search(r"D       =\s+([-+0-9.eEnaif]+)", txt).group(1)),
                      float(re.search(r"p value =\s+([-+0-9.eEnaif]+)", txt).group(1)))
    assert "-- Approximate Kolmogorov-Smirnov test --" in v_ours
    assert ks(v_ours) == ks(v_ref)


@pytest.mark.skipif(not os.path.exists(REF_CLI), reason="oracle/_ref/PROPhet not built")
@pytest.mark.parametrize("algorithm", ["bfgs", "steepest_descent_with_line_opt"])
def test_cli_line_search_training_matches_reference_cli(algorithm, tmp_path):
    """`algorithm = bfgs` / `steepest_descent_with_line_opt` (Optimizer.cpp:457-507, 607-869): gradients from the GPU,
    line search and L-BFGS recursion on the host; the printed table (one line per finished line search) and the saved
    potential must follow the reference's."""
    assert os.path.exists(hostapi.CLI_PATH), "PROPhet_b200 not built"
    cfg = synth.CONFIGS["C2"](6)
    ours = tmp_path / "ours"; ref = tmp_path / "ref"
    for d in (ours, ref):
        d.mkdir()
        _write_case(str(d), cfg, 6, [0, 0, 0, 0, 0, 0], [6, 5], seed=12, niter=8,
                    extra=("algorithm = " + algorithm, "line_min_epsilon = 1e-3"))
    out_ref = _run(REF_CLI, ["-in", "input_file", "-train"], str(ref))
    out_ours = _run(hostapi.CLI_PATH, ["-in", "input_file", "-train"], str(ours))
    t_ref, t_ours = _iteration_table(out_ref), _iteration_table(out_ours)
    assert len(t_ref) == 8 and len(t_ours) == 8, (out_ref[-1500:], out_ours[-1500:])
    assert np.allclose(np.array(t_ours), np.array(t_ref), rtol=5e-6, atol=0), (t_ours, t_ref)
    a = ckpt.read_checkpoint(str(ours / "trained_Si"))
    b = ckpt.read_checkpoint(str(ref / "trained_Si"))
    assert rel_err(a["params"], b["params"]) < 1e-7


@pytest.mark.skipif(not os.path.exists(REF_CLI), reason="oracle/_ref/PROPhet not built")
@pytest.mark.parametrize("extra", [("precondition_output = true",), ("sgd = true", "seed = 7"), ("regularization = 1e-3",),
                                   ("precondition_output = true", "regularization = 5e-4", "sgd = true", "seed = 3")])
def test_cli_training_options_match_reference_cli(extra, tmp_path):
    """The training options that change the arithmetic around the GPU epoch: output preconditioning
    (Potential.cpp:461-495), the shuffled visiting order of `sgd` (Potential.cpp:632-634: a full-batch pass in a different
    order, same libc rand() stream) and the L2 term of the optimizer (Optimizer.cpp:897-905).  Same input_file,
    PROPhet.xml and checkpoint_in for both programs; the training table and the saved potentials must agree."""
    assert os.path.exists(hostapi.CLI_PATH), "PROPhet_b200 not built"
    # `sgd` is compared on a single-element system: with several elements the reference's shuffled order trips over its
    # own accumulator reset (NN.cpp:629-646 resets a net's dOutput/dParameters when i_sys changes or when ITS cursor
    # wraps; a system visited last in one epoch and first in the next leaves the other elements' nets un-reset), which
    # corrupts the printed gradient norm of those epochs.  That stale sum is not reproduced.
    cfg = synth.CONFIGS["C2" if any(e.startswith("sgd") for e in extra) else "C3"](6)
    ours = tmp_path / "ours"; ref = tmp_path / "ref"
    for d in (ours, ref):
        d.mkdir()
        _write_case(str(d), cfg, 6, [0, 0, 1, 0, 0, 2], [7, 5], seed=15, niter=6, extra=extra)
    out_ref = _run(REF_CLI, ["-in", "input_file", "-train"], str(ref))
    out_ours = _run(hostapi.CLI_PATH, ["-in", "input_file", "-train"], str(ours))
    t_ref, t_ours = _iteration_table(out_ref), _iteration_table(out_ours)
    assert len(t_ref) == 6 and len(t_ours) == 6, (out_ref[-1500:], out_ours[-1500:])
    assert np.allclose(np.array(t_ours), np.array(t_ref), rtol=2e-6, atol=0), (t_ours, t_ref)
    for z in cfg.elements:
        sym = synth.Z_TO_SYMBOL[z]
        a = ckpt.read_checkpoint(str(ours / ("trained_" + sym)))
        b = ckpt.read_checkpoint(str(ref / ("trained_" + sym)))
        assert rel_err(a["params"], b["params"]) < 1e-9
        for k in ("out_mean", "out_scale"):
            if k in a and k in b:
                assert abs(a[k] - b[k]) <= 1e-11 * max(abs(b[k]), 1e-300)


@pytest.mark.parametrize("name", ["si_md", "tio2_md"])
def test_pair_style_nn_through_lammps_stand_in(name, tmp_path):
    """pair_style nn <Rcut> / pair_coeff <type> <file> / compute(), with checkpoint files in the reference's format."""
    c = MDCase(name)
    syms = [synth.Z_TO_SYMBOL[z] for z in c.elements]
    files = []
    for z in c.elements:
        p = str(tmp_path / ("pot_" + synth.Z_TO_SYMBOL[z]))
        ckpt.write_checkpoint(p, synth.Z_TO_SYMBOL[z], syms, {"G1": c.G1.tolist(), "G2": c.G2.tolist(), "G3": c.G3.tolist(), "G4": c.G4.tolist()},
                              c.d["in_mean.%d" % z], c.d["in_scale.%d" % z], c.out_mean, c.out_scale, c.sizes, c.d["params.%d" % z],
                              header_precision=17)
        files.append(p)
    # give the pair_coeff lines in reverse order: the type -> element map must come from the files, not the order
    types = list(range(1, len(c.elements) + 1))
    r = hostapi.pair_nn_run(c.rcut, types[::-1], files[::-1], c.x, c.types, c.nlocal, c.numneigh, c.neigh)
    assert rel_err(r["eatom"], c.d["eatom"]) < 1e-10
    assert abs(r["eng_vdwl"] - c.d["etotal"][0]) < 1e-10 * abs(c.d["etotal"][0])
    assert rel_err(r["f"], c.d["f"]) < 1e-10
    # init_style asked for the reference's neighbour settings (checked inside the harness), and a second PairNN built by
    # read_restart from the first one's write_restart record (pair_nn.cpp:322-398) computes the same energies
    assert r["restart_ok"]


def test_pair_style_nn_free_energy_shift(tmp_path):
    """Checkpoints carrying "FE:" lines: every per-atom energy becomes (E_i + sum of FE over locals and ghosts) / N
    (Potential.cpp:1159-1162); forces are untouched."""
    c = MDCase("tio2_md")
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "tio2_md_fe.npz"))
    syms = [synth.Z_TO_SYMBOL[z] for z in c.elements]
    files = []
    for k, z in enumerate(c.elements):
        p = str(tmp_path / ("pot_" + synth.Z_TO_SYMBOL[z]))
        ckpt.write_checkpoint(p, synth.Z_TO_SYMBOL[z], syms, {"G1": c.G1.tolist(), "G2": c.G2.tolist(), "G3": c.G3.tolist(), "G4": c.G4.tolist()},
                              c.d["in_mean.%d" % z], c.d["in_scale.%d" % z], c.out_mean, c.out_scale, c.sizes, c.d["params.%d" % z],
                              header_precision=17, free_energy=float(g["free_energy"][k]))
        files.append(p)
    types = list(range(1, len(c.elements) + 1))
    r = hostapi.pair_nn_run(c.rcut, types, files, c.x, c.types, c.nlocal, c.numneigh, c.neigh)
    assert rel_err(r["eatom"], g["eatom"]) < 1e-10
    assert abs(r["eng_vdwl"] - g["etotal"][0]) < 1e-10 * abs(g["etotal"][0])
    assert rel_err(r["f"], g["f"]) < 1e-10


@pytest.mark.skipif(not os.path.exists(REF_CLI), reason="oracle/_ref/PROPhet not built")
def test_cli_free_energy_training_and_validation(tmp_path):
    """`free_energy Ti <v> O <v>` (Setup.cpp free_energy key -> System.cpp:142-147): targets have the per-atom
    reference energies taken off before training, the saved potentials carry "FE:" lines, and -validate prints total
    energies with them added back (Potential.cpp:1001-1018,1046-1051)."""
    cfg = synth.CONFIGS["C3"](4)
    ours = tmp_path / "ours"; ref = tmp_path / "ref"
    fe = ("free_energy Ti -7.5 O -4.25",)
    for d in (ours, ref):
        d.mkdir()
        _write_case(str(d), cfg, 4, [0, 0, 0, 1], [6, 4], seed=14, niter=5, extra=fe)
    out_ref = _run(REF_CLI, ["-in", "input_file", "-train"], str(ref))
    out_ours = _run(hostapi.CLI_PATH, ["-in", "input_file", "-train"], str(ours))
    t_ref, t_ours = _iteration_table(out_ref), _iteration_table(out_ours)
    assert len(t_ref) == 5 and len(t_ours) == 5, (out_ref[-1500:], out_ours[-1500:])
    assert np.allclose(np.array(t_ours), np.array(t_ref), rtol=2e-6, atol=0), (t_ours, t_ref)
    for sym in ("O", "Ti"):
        a = open(str(ours / ("trained_" + sym))).read()
        b = open(str(ref / ("trained_" + sym))).read()
        assert [l for l in a.splitlines() if l.startswith("FE:")] == [l for l in b.splitlines() if l.startswith("FE:")] != []
    for d in (ours, ref):
        txt = open(str(d / "input_file")).read().replace("checkpoint_in = start", "checkpoint_in = trained")
        open(str(d / "input_file"), "w").write(txt)
    v_ref = _run(REF_CLI, ["-in", "input_file", "-validate"], str(ref))
    v_ours = _run(hostapi.CLI_PATH, ["-in", "input_file", "-validate"], str(ours))

    def table(txt):
        rows = []
        for line in txt.splitlines():
            tok = line.split()
            if len(tok) == 4 and tok[0].isdigit() and tok[3].isdigit():
                rows.append((int(tok[0]), float(tok[1]), float(tok[2]), int(tok[3])))
        return rows

    tr, to = table(v_ref), table(v_ours)
    assert len(tr) == 4 and len(to) == 4, (v_ref[-1500:], v_ours[-1500:])
    for a, b in zip(to, tr):
        assert a[0] == b[0] and a[3] == b[3]
        assert abs(a[1] - b[1]) <= 1e-8 * abs(b[1]) and abs(a[2] - b[2]) <= 1e-9 * abs(b[2])
    rms = lambda txt: float(re.search(r"RMS Error =\s+([-+0-9.eE]+)", txt).group(1))
    assert abs(rms(v_ours) - rms(v_ref)) <= 1e-5 * abs(rms(v_ref))


@pytest.mark.skipif(not os.path.exists(REF_CLI), reason="oracle/_ref/PROPhet not built")
def test_cli_optimize_follows_reference_monte_carlo(tmp_path):
    """`-optimize` (Potential::optimize_Gs, Potential.cpp:709-877): same libc rand() stream (seed key), so the perturbed
    rows, the retrained errors, the Metropolis decisions, logfile.out and G_parameters.out must all follow the
    reference.  The perturbed xi are non-integer: the descriptor refresh runs the binomial-series angular term.
    Inputs are vasprun.xml files (fractional coordinates): with Cartesian inputs the reference converts the positions
    to fractional AGAIN on every descriptor refresh (Structure.cpp:853-856 -> Normalize), which is not reproduced."""
    cfg = synth.CONFIGS["C2"](5)
    cfg.hidden = [5, 4]
    cfg.G2 = cfg.G2[:4]
    cfg.G4 = [cfg.G4[i] for i in (0, 3, 9, 14, 22, 31)]
    ours = tmp_path / "ours"; ref = tmp_path / "ref"
    extra = ("seed = 4242", "T0 = 1", "Tf = 0.85", "dT = 0.9", "Nsteps_per_T = 3", "step_size = 0.1")
    for d in (ours, ref):
        d.mkdir()
        data = ["data code=vasp"]
        for k in range(5):
            cell, cart, Z, target, _ = synth.make_structure(cfg, 40 + k)
            synth.write_vasprun(str(d / ("run%d" % k)), cell, cart @ np.linalg.inv(cell), Z, target)
            data.append("./run%d%s" % (k, "  val" if k == 4 else ""))
        net = synth.random_network(cfg, 14, 15)
        ckpt.write_checkpoint(str(d / "start_Si"), "Si", ["Si"], {"G1": cfg.G1, "G2": cfg.G2, "G3": cfg.G3, "G4": cfg.G4},
                              0.3 * np.abs(synth.normal(14, cfg.NG, 3)), 0.5 + np.abs(synth.normal(14, cfg.NG, 4)), 0.0, 1.0,
                              net["sizes"], net["params"], header_precision=17)
        synth.write_input_file(str(d / "input_file"), cfg, niterations=4, checkpoint_in="start", checkpoint_out="trained",
                               extra=extra, data=data)
    out_ref = _run(REF_CLI, ["-in", "input_file", "-optimize"], str(ref))
    out_ours = _run(hostapi.CLI_PATH, ["-in", "input_file", "-optimize"], str(ours))
    err = lambda txt: [float(m) for m in re.findall(r"^Error =\s+([-+0-9.eE]+)", txt, flags=re.M)]
    e_ref, e_ours = err(out_ref), err(out_ours)
    assert len(e_ref) == 6 and len(e_ours) == 6, (out_ref[-800:], out_ours[-800:])
    assert np.allclose(e_ours, e_ref, rtol=2e-5), (e_ours, e_ref)
    t_ref, t_ours = _iteration_table(out_ref), _iteration_table(out_ours)
    assert len(t_ref) == 24 and np.allclose(np.array(t_ours), np.array(t_ref), rtol=5e-6), (t_ours, t_ref)
    num = lambda path: [[float(x) for x in re.findall(r"[-+]?[0-9.]+(?:e[-+]?[0-9]+)?", l)] for l in open(path)
                        if l.strip() and not l.startswith(" -") and "accepted" not in l]
    assert np.allclose(np.array(num(str(ours / "logfile.out"))), np.array(num(str(ref / "logfile.out"))), rtol=1e-5)
    ga, gb = num(str(ours / "G_parameters.out")), num(str(ref / "G_parameters.out"))
    assert len(ga) == len(gb) == 10
    for ra, rb in zip(ga, gb):
        assert len(ra) == len(rb) and np.allclose(ra, rb, rtol=1e-5), (ra, rb)


@pytest.mark.skipif(not os.path.exists(REF_CLI), reason="oracle/_ref/PROPhet not built")
def test_cli_print_gvector_file_matches_reference(tmp_path):
    """print_gvector = true (Potential.cpp:114-136): both programs dump every atom's descriptors to 0.gvector."""
    cfg = synth.CONFIGS["C3"](3)
    ours = tmp_path / "ours"; ref = tmp_path / "ref"
    for d in (ours, ref):
        d.mkdir()
        _write_case(str(d), cfg, 3, [0, 0, 1], [5, 4], seed=16, niter=1, extra=("print_gvector = true",))
    _run(REF_CLI, ["-in", "input_file", "-train"], str(ref))
    _run(hostapi.CLI_PATH, ["-in", "input_file", "-train"], str(ours))

    def load(path):
        blocks, cur = [], []
        for line in open(path):
            if line.strip() == "":
                if cur:
                    blocks.append(np.array(cur)); cur = []
            else:
                cur.append([float(v) for v in line.strip().rstrip(",").split(",")])
        if cur:
            blocks.append(np.array(cur))
        return blocks

    a, b = load(str(ours / "0.gvector")), load(str(ref / "0.gvector"))
    assert len(a) == len(b) == 3
    for x, y in zip(a, b):
        assert x.shape == y.shape == (cfg.natoms, cfg.NG)
        assert np.allclose(x, y, rtol=2e-5, atol=1e-12)
