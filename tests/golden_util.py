"""Helpers that turn tests/golden/*.npz (reference outputs) into ready-to-use test inputs."""
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

TRAIN_CASES = ["si_train", "tio2_train", "abo4_train", "variants_train", "tutorial_c1", "batio3_doc"]
MD_CASES = ["si_md", "tio2_md", "variants_md"]


def load(name):
    return dict(np.load(os.path.join(GOLDEN_DIR, name + ".npz")))


class TrainCase:
    """A training golden: structures, rows, reference G / neighbour lists / per-iteration SSE, gradient, x."""

    def __init__(self, name):
        d = load(name)
        self.name = name
        self.d = d
        self.n = int(d["n_systems"][0])
        self.elements = [int(z) for z in d["atom_types"]]
        self.rcut = float(d["Rcut"][0])
        self.G1, self.G2, self.G3, self.G4 = d["G1p"], d["G2p"], d["G3p"], d["G4p"]
        self.NG = int(d["NG"][0])
        zmap = {z: i for i, z in enumerate(self.elements)}
        self.cells = [d["cell.%d" % s] for s in range(self.n)]
        self.frac = [d["frac_pos.%d" % s] for s in range(self.n)]
        self.raw = [d["raw_pos.%d" % s] for s in range(self.n)]
        self.raw_is_cart = [bool(d["raw_is_cartesian.%d" % s][0]) for s in range(self.n)]
        self.Z = [d["Z.%d" % s] for s in range(self.n)]
        self.elem = [np.array([zmap[int(z)] for z in Zs], dtype=np.int32) for Zs in self.Z]
        self.targets = np.array([d["target.%d" % s][0] for s in range(self.n)])
        self.tags = np.array([int(d["tag.%d" % s][0]) for s in range(self.n)], dtype=np.int32)
        self.G = [d["G.%d" % s] for s in range(self.n)]            # [NG][natom]
        self.natoms = np.array([len(z) for z in self.Z], dtype=np.int32)
        self.atom_offset = np.zeros(self.n + 1, dtype=np.int32)
        self.atom_offset[1:] = np.cumsum(self.natoms)
        self.n_iter = len([k for k in d if k.startswith("SSE.")])
        self.hidden = [int(h) for h in d["hidden"]]
        self.sizes = [self.NG] + self.hidden + [1]
        self.P_net = sum((a + 1) * b for a, b in zip(self.sizes[:-1], self.sizes[1:]))
        self.out_mean = float(d["output_mean"][0]) if "output_mean" in d else 0.0
        self.out_scale = float(d["output_variance"][0]) if "output_variance" in d else 1.0

    def G_atom_major(self):
        return np.concatenate([g.T for g in self.G], axis=0).copy()

    def elem_flat(self):
        return np.concatenate(self.elem)

    def node_transfer(self, z):
        key = "node_transfer.%d" % z
        if key in self.d:
            return [str(x) for x in self.d[key]]
        n_nodes = sum(self.sizes[1:])
        return [str(self.d["transfer"])] * (n_nodes - 1) + ["linear"]

    def mean(self, z):
        return self.d["input_mean.%d" % z]

    def scale(self, z):
        return self.d["input_scale.%d" % z]

    def params_at(self, it):
        """Per-element parameter vectors at the start of iteration ``it`` (slices of x_before, ascending Z)."""
        x = self.d["x_before.%d" % it]
        return [x[e * self.P_net:(e + 1) * self.P_net] for e in range(len(self.elements))]


class MDCase:
    def __init__(self, name):
        d = load(name)
        self.name = name
        self.d = d
        self.x = d["x"]
        self.types = d["types"]
        self.elem = (d["types"] - 1).astype(np.int32)      # LAMMPS type t -> element index t-1 (ascending Z)
        self.nlocal = int(d["nlocal"])
        self.numneigh = d["numneigh"]
        self.offsets = d["neigh_offsets"].astype(np.int64)
        self.neigh = d["neigh_flat"].astype(np.int32)
        self.rcut = float(d["Rcut"])
        self.elements = [int(z) for z in d["elements"]]
        self.G1, self.G2, self.G3, self.G4 = d["G1p"], d["G2p"], d["G3p"], d["G4p"]
        self.NG = int(d["NG"][0])
        self.hidden = [int(h) for h in d["hidden"]]
        self.sizes = [self.NG] + self.hidden + [1]
        self.out_mean = float(d["out_mean"])
        self.out_scale = float(d["out_scale"])

    def node_transfer(self, z):
        key = "node_transfer.%d" % z
        if key in self.d:
            return [str(x) for x in self.d[key]]
        n_nodes = sum(self.sizes[1:])
        return [str(self.d["transfer"])] * (n_nodes - 1) + ["linear"]


def rel_err(a, b):
    """max|a-b| relative to max|b| (relative to the max-abs of each array)."""
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    den = np.max(np.abs(b)) if b.size else 1.0
    if den == 0:
        den = 1.0
    return float(np.max(np.abs(a - b)) / den) if b.size else 0.0
