#!/usr/bin/env python
"""Generate the committed golden fixtures (tests/golden/*.npz) by running the UNMODIFIED reference.

Runs only in the build container (needs /root/reference and oracle/_ref built by oracle/build_ref.sh);
the fixtures it writes travel with the repo, the reference does not.  Every array in a fixture is an
output of the reference's own code (via oracle/ref_driver.cpp) or an input that was fed to it.

    python tests/golden/make_goldens.py            # regenerate everything
    python tests/golden/make_goldens.py si_train   # one case
"""
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
sys.path.insert(0, ROOT)
from oracle import refdump  # noqa: E402
from prophet_b200 import ckpt, synth  # noqa: E402

REF = os.environ.get("PROPHET_REFERENCE", "/root/reference")
DRIVER = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
OUT = os.path.dirname(os.path.abspath(__file__))


def run_driver(args, cwd):
    r = subprocess.run([DRIVER] + args, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("ref_driver failed in %s:\n%s" % (cwd, r.stdout[-4000:]))
    return r.stdout


def rows_of(cfg):
    return {"G1": cfg.G1, "G2": cfg.G2, "G3": cfg.G3, "G4": cfg.G4}


def write_checkpoints(cwd, cfg, base, seed, node_transfer=None, means=None, scales=None):
    """Random nets for every element of cfg, written in the reference's format; returns per-Z dict."""
    syms = [synth.Z_TO_SYMBOL[z] for z in cfg.elements]
    nets = {}
    for z in cfg.elements:
        net = synth.random_network(cfg, z, seed)
        m = np.zeros(cfg.NG) if means is None else means[z]
        s = np.ones(cfg.NG) if scales is None else scales[z]
        ckpt.write_checkpoint(os.path.join(cwd, "%s_%s" % (base, synth.Z_TO_SYMBOL[z])), synth.Z_TO_SYMBOL[z], syms,
                              rows_of(cfg), m, s, 0.0, 1.0, net["sizes"], net["params"], transfer=cfg.transfer,
                              node_transfer=None if node_transfer is None else node_transfer[z],
                              header_precision=17, weight_precision=18)
        nets[z] = net
    return nets


def train_case(name, cfg, structures, n_iter, seed, *, precondition_from_checkpoint=False, extra=(), node_transfer=None,
               with_checkpoint=True, keep=None):
    """PROPhet.xml + input_file (+ checkpoints) -> ref_driver train -> npz."""
    cwd = tempfile.mkdtemp(prefix="golden_" + name + "_")
    try:
        synth.write_prophet_xml(os.path.join(cwd, "PROPhet.xml"), structures)
        arrays = {}
        if with_checkpoint:
            # a checkpoint loaded by the reference carries its own means/scales and so disables input
            # preconditioning (NN.h:100-103).  To exercise Precondition() we therefore run the driver twice:
            # once without a checkpoint to get the reference's own means/scales, then with a checkpoint that
            # carries them.
            means = scales = None
            if not precondition_from_checkpoint:
                synth.write_input_file(os.path.join(cwd, "input_pre"), cfg, niterations=1, extra=("seed = 7",) + tuple(extra))
                run_driver(["train", "input_pre", "pre.bin", "1"], cwd)
                pre = refdump.load(os.path.join(cwd, "pre.bin"))
                means = {z: pre["input_mean.%d" % z] for z in cfg.elements}
                scales = {z: pre["input_scale.%d" % z] for z in cfg.elements}
                for z in cfg.elements:
                    arrays["precond_mean.%d" % z] = means[z]
                    arrays["precond_scale.%d" % z] = scales[z]
            write_checkpoints(cwd, cfg, "start", seed, node_transfer=node_transfer, means=means, scales=scales)
            synth.write_input_file(os.path.join(cwd, "input_file"), cfg, niterations=n_iter + 1, checkpoint_in="start",
                                   extra=extra)
        else:
            synth.write_input_file(os.path.join(cwd, "input_file"), cfg, niterations=n_iter + 1,
                                   extra=("seed = 7",) + tuple(extra))
        run_driver(["train", "input_file", "out.bin", str(n_iter)], cwd)
        d = refdump.load(os.path.join(cwd, "out.bin"))
        arrays.update(d)
        if keep is not None:      # optimizer-only fixtures: drop the per-structure dumps
            arrays = {k: v for k, v in arrays.items() if k.split(".")[0] in keep}
        arrays["hidden"] = np.array(cfg.hidden, dtype=np.int32)
        arrays["transfer"] = np.array(cfg.transfer)
        if node_transfer is not None:
            for z, names in node_transfer.items():
                arrays["node_transfer.%d" % z] = np.array(names)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **arrays)
        print("wrote %s.npz (%d arrays)" % (name, len(arrays)))
    finally:
        shutil.rmtree(cwd, ignore_errors=True)


def md_case(name, cfg, structure, seed, *, skin=1.0, node_transfer=None, free_energy=None, keep=None):
    """Ghost-expanded cell + full neighbour list -> ref_driver md -> npz."""
    cell, cart, Z, _, _ = structure
    cwd = tempfile.mkdtemp(prefix="golden_" + name + "_")
    try:
        # element -> LAMMPS type 1..ntypes in ascending Z (type ids stand in for atomic numbers, Structure.cpp:1406)
        type_of = {z: i + 1 for i, z in enumerate(cfg.elements)}
        syms = [synth.Z_TO_SYMBOL[z] for z in cfg.elements]
        rng_means = {z: 0.1 * synth.normal(seed + z, cfg.NG, stream=3) for z in cfg.elements}
        rng_scales = {z: 0.5 + np.abs(synth.normal(seed + z, cfg.NG, stream=4)) for z in cfg.elements}
        nets = {}
        for z in cfg.elements:
            net = synth.random_network(cfg, z, seed)
            ckpt.write_checkpoint(os.path.join(cwd, "pot_%s" % synth.Z_TO_SYMBOL[z]), synth.Z_TO_SYMBOL[z], syms,
                                  rows_of(cfg), rng_means[z], rng_scales[z], -1.25, 1.5, net["sizes"], net["params"],
                                  transfer=cfg.transfer,
                                  node_transfer=None if node_transfer is None else node_transfer[z],
                                  header_precision=17, weight_precision=18,
                                  free_energy=None if free_energy is None else free_energy[z])
            nets[z] = net
        x, Zall, nlocal = synth.expand_ghosts(cell, cart, Z, cfg.rcut + skin)
        numneigh, offsets, flat = synth.brute_full_neighbor_list(x, nlocal, cfg.rcut + skin)
        types = np.array([type_of[int(z)] for z in Zall], dtype=np.int32)
        with open(os.path.join(cwd, "case.txt"), "w") as f:
            f.write("Rcut %r\n" % float(cfg.rcut))
            f.write("ntypes %d\n" % len(cfg.elements))
            for z in cfg.elements:
                f.write("coeff %d pot_%s\n" % (type_of[z], synth.Z_TO_SYMBOL[z]))
            f.write("atoms %d %d\n" % (nlocal, len(x) - nlocal))
            for t, r in zip(types, x):
                f.write("%d %r %r %r\n" % (t, float(r[0]), float(r[1]), float(r[2])))
            for i in range(nlocal):
                f.write("neigh %d %d %s\n" % (i, numneigh[i], " ".join(str(j) for j in flat[offsets[i]:offsets[i + 1]])))
        run_driver(["md", "case.txt", "out.bin"], cwd)
        d = refdump.load(os.path.join(cwd, "out.bin"))
        arrays = dict(d)
        arrays.update(x=x, types=types, nlocal=np.array(nlocal), numneigh=numneigh, neigh_offsets=offsets,
                      neigh_flat=flat, Rcut=np.array(cfg.rcut), elements=np.array(cfg.elements, dtype=np.int32),
                      hidden=np.array(cfg.hidden, dtype=np.int32), transfer=np.array(cfg.transfer),
                      out_mean=np.array(-1.25), out_scale=np.array(1.5),
                      G1p=np.array(cfg.G1, dtype=np.float64).reshape(-1, 2), G2p=np.array(cfg.G2, dtype=np.float64).reshape(-1, 4),
                      G3p=np.array(cfg.G3, dtype=np.float64).reshape(-1, 5), G4p=np.array(cfg.G4, dtype=np.float64).reshape(-1, 5))
        for z in cfg.elements:
            arrays["params.%d" % z] = nets[z]["params"]
            arrays["in_mean.%d" % z] = rng_means[z]
            arrays["in_scale.%d" % z] = rng_scales[z]
            if node_transfer is not None:
                arrays["node_transfer.%d" % z] = np.array(node_transfer[z])
        if keep is not None:
            arrays = {k: v for k, v in arrays.items() if k in keep}
        if free_energy is not None:
            arrays["free_energy"] = np.array([free_energy[z] for z in cfg.elements])
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **arrays)
        print("wrote %s.npz (%d arrays, %d local + %d ghost atoms)" % (name, len(arrays), nlocal, len(x) - nlocal))
    finally:
        shutil.rmtree(cwd, ignore_errors=True)


def retag(structs, tags):
    return [(c, x, z, t, tag) for (c, x, z, t, _), tag in zip(structs, tags)]


# ----------------------------------------------------------------------------------------------
def variants_config():
    """Small TiO2 cell with every descriptor / cutoff variant the reference implements: G1, G3, fc type 1,
    non-integer xi (binomial series), per-row Rc, and a sheared (triclinic) cell."""
    lat = synth.lattice_tio2((1, 1, 2))
    cell = lat[0].copy()
    cell[1, 0] = 0.7    # shear b towards a
    cell[2, 1] = -0.4   # and c towards -b
    frac = lat[1] @ np.linalg.inv(lat[0])
    cfg = synth.Config("variants", 9, (cell, frac @ cell, lat[2]), 3,
                       G2=[[4.5, 0.0, 0.8, 0.0], [4.0, 1.0, 0.3, 1.5], [4.5, 0.0, 2.0, 2.0]],
                       G4=[[4.5, 0.0, 0.05, 1.0, 1.0], [4.5, 0.0, 0.05, 2.5, -1.0], [4.5, 0.0, 0.2, 3.0, 1.0],
                           [3.8, 1.0, 0.2, 1.7, 1.0]],
                       hidden=[6, 5], rcut=4.5, sigma=0.05,
                       G1=[[4.5, 0.0], [3.5, 1.0]],
                       G3=[[4.5, 0.0, 0.02, 1.0, 1.0], [4.5, 0.0, 0.05, 2.0, -1.0], [4.0, 1.0, 0.03, 1.5, 1.0]])
    return cfg


ALL_TRANSFERS = ["tanh", "tanh_spline", "tanh_crazy", "cubic", "pulse", "sin", "periodic_tanh", "relu", "linear"]


def case_si_train():
    cfg = synth.config_c2(3)
    st = retag([synth.make_structure(cfg, s) for s in range(3)], [0, 1, 0])
    train_case("si_train", cfg, st, 4, seed=21)


def case_tio2_train():
    cfg = synth.config_c3(3)
    st = retag([synth.make_structure(cfg, s) for s in range(3)], [0, 0, 1])
    train_case("tio2_train", cfg, st, 3, seed=31)


def case_abo4_train():
    cfg = synth.config_c4(2)      # BASELINE config 4 as named: 120 descriptors, three nets 120-60-60-60-1 (P = 43 923)
    st = retag([synth.make_structure(cfg, s) for s in range(2)], [0, 0])
    train_case("abo4_train", cfg, st, 2, seed=41)


def case_variants_train():
    cfg = variants_config()
    st = retag([synth.make_structure(cfg, s) for s in range(3)], [0, 0, 2])
    # Functional_params::print crashes on G1 rows (Functional_params.cpp:155-156) but reading them is fine.
    nt = {}
    for z in cfg.elements:
        names = []
        k = 0
        for h in cfg.hidden:
            for _ in range(h):
                names.append(ALL_TRANSFERS[k % len(ALL_TRANSFERS)])
                k += 1
        nt[z] = names + ["linear"]
    train_case("variants_train", cfg, st, 2, seed=51, node_transfer=nt)


OPT_KEEP = ("x_before", "x_after", "SSE", "vSSE", "nval", "gradient", "Ntrain")


def _opt_case(name, algorithm_lines, n_calls, seed, tiny=False):
    """Optimizer-only fixture: every Optimizer::update_network call of a short reference training run (its inputs
    SSE/gradient and the parameter vector it hands back), for the line-search algorithms."""
    cfg = synth.config_c2(4)
    cfg.hidden = [2] if tiny else [5, 4]
    cfg.G2 = cfg.G2[:2] if tiny else cfg.G2[:6]
    cfg.G4 = [] if tiny else cfg.G4[:6]
    st = retag([synth.make_structure(cfg, 70 + s) for s in range(4)], [0, 0, 0, 0 if tiny else 1])
    train_case(name, cfg, st, n_calls, seed=seed, precondition_from_checkpoint=True, extra=algorithm_lines, keep=OPT_KEEP)


def case_opt_bfgs():
    _opt_case("opt_bfgs", ("algorithm = bfgs", "line_min_epsilon = 1e-3"), 60, seed=61)


def case_opt_sd_line():
    _opt_case("opt_sd_line", ("algorithm = steepest_descent_with_line_opt", "line_min_epsilon = 1e-3"), 40, seed=62)


def case_opt_default_switch():
    # a 9-parameter net: the "default" schedule leaves Rprop for steepest descent at iteration 134 (2 x 67) and
    # steepest descent for L-BFGS at iteration 150; the reference's L-BFGS then stalls on the flat bottom and
    # turns to NaN at call 211 (y.s = 0 -> rho = inf) -- kept in the fixture on purpose.
    _opt_case("opt_default_switch", ("algorithm = default", "debug = 0.5", "line_min_epsilon = 1e-3"), 215, seed=63,
              tiny=True)


def case_variants_md():
    # orthorhombic version of the variants rows (ghost expansion helper wants a rectangular box); no G3 quirk-free
    # shortcut: G3 rows are present, so the "Rjk > Rcut skips G4" branch (Structure.cpp:1129-1132) is exercised.
    # G4 rows: one Rc / fc type and eta-major order with >= 2 rows per eta -- the only patterns for which the
    # reference's row-to-row memoisation in Get_Forces (Structure.cpp:656-698) yields the intended derivative
    # (SURVEY.md appendix C #4); fc type 1 and non-integer xi are still covered.
    v = variants_config()
    g4 = [[4.5, 1.0, 0.05, 1.0, 1.0], [4.5, 1.0, 0.05, 2.5, -1.0], [4.5, 1.0, 0.2, 3.0, 1.0], [4.5, 1.0, 0.2, 1.7, 1.0]]
    cfg = synth.Config("variants_md", 10, synth.lattice_tio2((2, 2, 3)), 1, G2=v.G2, G4=g4, hidden=[6, 5], rcut=4.5,
                       sigma=0.05, G1=v.G1, G3=v.G3)
    nt = {}
    for z in cfg.elements:
        names = [ALL_TRANSFERS[k % len(ALL_TRANSFERS)] for k in range(sum(cfg.hidden))]
        nt[z] = names + ["linear"]
    md_case("variants_md", cfg, synth.make_structure(cfg, 0), seed=52, node_transfer=nt)


def case_si_md():
    cfg = synth.config_c2(1)
    md_case("si_md", cfg, synth.make_structure(cfg, 5), seed=61)


def case_tio2_md():
    cfg = synth.config_c3(1)
    cfg.lattice = synth.lattice_tio2((2, 2, 3))
    md_case("tio2_md", cfg, synth.make_structure(cfg, 6), seed=71)


def case_tio2_md_fe():
    """tio2_md again with "FE:" lines in both checkpoints: only the energies change (Potential.cpp:1159-1162), so only
    they are kept; the inputs are tio2_md.npz's."""
    cfg = synth.config_c3(1)
    cfg.lattice = synth.lattice_tio2((2, 2, 3))
    md_case("tio2_md_fe", cfg, synth.make_structure(cfg, 6), seed=71, free_energy={8: -4.25, 22: -7.5},
            keep=("eatom", "etotal", "f"))


def case_tutorial_c1():
    """First 4 shipped tutorial structures through the reference's VASP reader, auto-generated rows
    (Nradial=5 Nangular=4, Structure.cpp:818-851), net 9-20-1 tanh, reference-initialised weights (seed 135492)."""
    cwd = tempfile.mkdtemp(prefix="golden_c1_")
    try:
        tut = os.path.join(REF, "doc", "tutorial")
        os.symlink(os.path.join(tut, "data"), os.path.join(cwd, "data"))
        with open(os.path.join(tut, "training_data")) as f:
            dirs = [ln.split()[0] for ln in f if ln.strip().startswith("./data/")]
        with open(os.path.join(cwd, "input_file"), "w") as f:
            f.write("network_type = nn\nhidden = 20\ninput = structure\noutput = energy\ntransfer = tanh\n"
                    "Nradial=5\nNangular=4\nRcut=4\nalgorithm=rprop\nNiterations = 7\nthreshold = 0\nNprint = 1\n"
                    "checkpoint_out = potential\nNcheckpoint = 1000000\nseed = 135492\ndata code=vasp\n")
            for k, dname in enumerate(dirs[:4]):
                f.write("%s%s\n" % (dname, "  val" if k == 3 else ""))
        run_driver(["train", "input_file", "out.bin", "6"], cwd)
        d = refdump.load(os.path.join(cwd, "out.bin"))
        d["hidden"] = np.array([20], dtype=np.int32)
        d["transfer"] = np.array("tanh")
        np.savez_compressed(os.path.join(OUT, "tutorial_c1.npz"), **d)
        print("wrote tutorial_c1.npz (%d arrays)" % len(d))
    finally:
        shutil.rmtree(cwd, ignore_errors=True)


def case_batio3_doc():
    """First 4 systems of the shipped ternary fixture doc/PROPhet.xml through the code=prophet reader."""
    cwd = tempfile.mkdtemp(prefix="golden_batio3_")
    try:
        with open(os.path.join(REF, "doc", "PROPhet.xml")) as f:
            text = f.read()
        head, rest = text.split("<systems>", 1)
        systems = rest.split("</system>")
        keep = "</system>".join(systems[:4]) + "</system>\n  </systems>\n</PROPhet>\n"
        head = head.replace("<nsystem>100</nsystem>", "<nsystem>4</nsystem>")
        with open(os.path.join(cwd, "PROPhet.xml"), "w") as f:
            f.write(head + "<systems>" + keep)
        with open(os.path.join(cwd, "input_file"), "w") as f:
            f.write("network_type = nn\nhidden = 10 10\ninput = structure\noutput = energy\ntransfer = tanh\n"
                    "Nradial=3\nNangular=4\nRcut=5.0\nalgorithm=rprop\nNiterations = 5\nthreshold = 0\nNprint = 1\n"
                    "checkpoint_out = potential\nNcheckpoint = 1000000\nseed = 99\ndata code=prophet\n")
        run_driver(["train", "input_file", "out.bin", "4"], cwd)
        d = refdump.load(os.path.join(cwd, "out.bin"))
        d["hidden"] = np.array([10, 10], dtype=np.int32)
        d["transfer"] = np.array("tanh")
        np.savez_compressed(os.path.join(OUT, "batio3_doc.npz"), **d)
        print("wrote batio3_doc.npz (%d arrays)" % len(d))
    finally:
        shutil.rmtree(cwd, ignore_errors=True)


CASES = {
    "si_train": case_si_train, "tio2_train": case_tio2_train, "abo4_train": case_abo4_train,
    "variants_train": case_variants_train, "si_md": case_si_md, "tio2_md": case_tio2_md,
    "variants_md": case_variants_md, "tutorial_c1": case_tutorial_c1, "batio3_doc": case_batio3_doc,
}



def case_checkpoint_files():
    """Checkpoint files written by the reference's own Potential::save (3 Rprop iterations on two small TiO2
    structures, nets 52-5-4-1, tanh_spline default transfer): byte-level fixtures for the host layer's reader/writer."""
    cfg = synth.config_c3(2)
    cfg.hidden = [5, 4]
    cfg.transfer = "tanh_spline"
    cwd = tempfile.mkdtemp(prefix="golden_ckpt_")
    try:
        st = [synth.make_structure(cfg, s) for s in range(2)]
        synth.write_prophet_xml(os.path.join(cwd, "PROPhet.xml"), st)
        synth.write_input_file(os.path.join(cwd, "input_file"), cfg, niterations=3, extra=("seed = 11",))
        r = subprocess.run([os.path.join(ROOT, "oracle", "_ref", "PROPhet"), "-in", "input_file", "-train"], cwd=cwd,
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        assert r.returncode == 0, r.stdout[-2000:]
        for sym in ("O", "Ti"):
            shutil.copy(os.path.join(cwd, "potential_" + sym), os.path.join(OUT, "ref_checkpoint_" + sym + ".txt"))
        print("wrote ref_checkpoint_O.txt, ref_checkpoint_Ti.txt")
    finally:
        shutil.rmtree(cwd, ignore_errors=True)


CASES["checkpoint_files"] = case_checkpoint_files


def case_ks():
    """Analysis::KS on a few seeded sample pairs (N = 8 -> p not evaluated, N = 64 and 200 -> series evaluated,
    identical samples -> D = 0, p = 1)."""
    arrays = {}
    cwd = tempfile.mkdtemp(prefix="golden_ks_")
    try:
        for k, (n, shift, scale) in enumerate([(8, 0.3, 1.0), (64, 0.05, 1.1), (200, 0.4, 0.7), (40, 0.0, 1.0)]):
            a = synth.normal(900 + k, n, 1)
            b = a.copy() if (shift == 0.0 and scale == 1.0) else shift + scale * synth.normal(900 + k, n, 2)
            with open(os.path.join(cwd, "s.txt"), "w") as f:
                f.write("%d\n%s\n%d\n%s\n" % (n, " ".join(repr(float(v)) for v in a), n, " ".join(repr(float(v)) for v in b)))
            run_driver(["ks", "s.txt", "ks.bin"], cwd)
            d = refdump.load(os.path.join(cwd, "ks.bin"))
            arrays["a.%d" % k], arrays["b.%d" % k] = a, b
            arrays["D.%d" % k], arrays["p.%d" % k] = d["D"], d["p"]
        np.savez_compressed(os.path.join(OUT, "ks.npz"), **arrays)
        print("wrote ks.npz", [(float(arrays["D.%d" % k][0]), float(arrays["p.%d" % k][0])) for k in range(4)])
    finally:
        shutil.rmtree(cwd, ignore_errors=True)


def case_opt_perturbation():
    """The perturbation ("global optimizer") scheme of Optimizer::update_network (Optimizer.cpp:915-938, 1008-1043) on a
    quadratic objective, P = 9, Rprop, lambda_max = 1.5 (the accessor truncates it to 1), threshold 5e-3, seed 77:
    the run turns the perturbation on twice, detects a transition, continues, finds none, and stops at call 12560."""
    cwd = tempfile.mkdtemp(prefix="golden_optq_")
    try:
        spec = dict(P=9, ncalls=40000, algorithm="rprop", lambda_max=1.5, ncycles=1, threshold=5e-3, seed=77)
        with open(os.path.join(cwd, "spec.txt"), "w") as f:
            f.write("%(P)d %(ncalls)d %(algorithm)s %(lambda_max)r %(ncycles)d %(threshold)r %(seed)d\n" % spec)
        run_driver(["optq", "spec.txt", "out.bin"], cwd)
        d = dict(refdump.load(os.path.join(cwd, "out.bin")))
        for k, v in spec.items():
            d["spec_" + k] = np.array(v)
        np.savez_compressed(os.path.join(OUT, "opt_perturbation.npz"), **d)
        print("wrote opt_perturbation.npz: converged at call %d" % int(d["converged_at"][0]))
    finally:
        shutil.rmtree(cwd, ignore_errors=True)


CASES["opt_perturbation"] = case_opt_perturbation
CASES["ks"] = case_ks
CASES["tio2_md_fe"] = case_tio2_md_fe
CASES["opt_bfgs"] = case_opt_bfgs
CASES["opt_sd_line"] = case_opt_sd_line
CASES["opt_default_switch"] = case_opt_default_switch

if __name__ == "__main__":
    if not os.path.exists(DRIVER):
        sys.exit("oracle/_ref/ref_driver missing: run oracle/build_ref.sh first (needs %s)" % REF)
    names = sys.argv[1:] or list(CASES)
    for n in names:
        CASES[n]()
