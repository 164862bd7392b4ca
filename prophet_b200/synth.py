"""Deterministic synthetic structure generator for the BASELINE.json configs.

Harness code (tests, bench, golden generation) -- not part of the compute path.

Structure ``s`` of config ``c`` uses seed ``c*10**9 + s``; positions are ideal lattice sites plus
N(0, sigma=0.08 A) per Cartesian coordinate, wrapped into the cell; targets are
``e0*Natom + N(0, 0.5)`` eV; every 20th structure is tagged ``val``, the rest ``train``.
The RNG is counter based (splitmix64 -> Box-Muller), so any structure can be generated
independently of the others and identically by any tool that implements the same few lines.

The same structures are written as ``PROPhet.xml`` (the reference's ``code=prophet`` reader,
CUSTOM.cpp:81-160) for the oracle and handed to the GPU path as flat arrays.
"""
from __future__ import annotations

import dataclasses
import os
from typing import Dict, List, Sequence, Tuple

import numpy as np

TAG_TRAIN, TAG_VAL, TAG_TEST = 0, 1, 2
TAG_NAMES = {TAG_TRAIN: "train", TAG_VAL: "val", TAG_TEST: "test"}

SYMBOL_TO_Z = {"C": 6, "O": 8, "Si": 14, "Ca": 20, "Ti": 22, "Ba": 56, "W": 74}
Z_TO_SYMBOL = {v: k for k, v in SYMBOL_TO_Z.items()}

_MASK = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x: np.ndarray) -> np.ndarray:
    """One splitmix64 output per 64-bit input counter (vectorised, wrap-around arithmetic)."""
    with np.errstate(over="ignore"):
        z = (x.astype(np.uint64) + np.uint64(0x9E3779B97F4A7C15)) & _MASK
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _MASK
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _MASK
        return z ^ (z >> np.uint64(31))


def uniform01(seed: int, counters: np.ndarray) -> np.ndarray:
    """Uniform doubles in (0,1): top 53 bits of splitmix64(seed*2^32 + counter), offset by half an ulp."""
    with np.errstate(over="ignore"):
        base = (np.uint64(seed) * np.uint64(0x100000000)) & _MASK
        bits = splitmix64(base + counters.astype(np.uint64))
    return ((bits >> np.uint64(11)).astype(np.float64) + 0.5) * (1.0 / 9007199254740992.0)


def normal(seed: int, n: int, stream: int = 0) -> np.ndarray:
    """n standard normals from counters [stream*2^20, ...) by Box-Muller (cos branch only)."""
    c = np.arange(n, dtype=np.uint64) * np.uint64(2) + np.uint64(stream) * np.uint64(1 << 20)
    u1 = uniform01(seed, c)
    u2 = uniform01(seed, c + np.uint64(1))
    return np.sqrt(-2.0 * np.log(u1)) * np.cos(2.0 * np.pi * u2)


# ----------------------------------------------------------------------------------------------
# lattices
def _diamond_basis() -> np.ndarray:
    fcc = np.array([[0, 0, 0], [0, .5, .5], [.5, 0, .5], [.5, .5, 0]], dtype=np.float64)
    return np.concatenate([fcc, fcc + 0.25])


def _supercell(cell: np.ndarray, basis_frac: np.ndarray, species: Sequence[str], reps: Tuple[int, int, int]):
    """Cartesian positions of an nx*ny*nz supercell; cell rows are the lattice vectors a,b,c."""
    nx, ny, nz = reps
    shifts = np.array([[i, j, k] for i in range(nx) for j in range(ny) for k in range(nz)], dtype=np.float64)
    frac = (shifts[:, None, :] + basis_frac[None, :, :]).reshape(-1, 3)
    cart = frac @ cell
    sp = list(species) * len(shifts)
    big = cell * np.array(reps, dtype=np.float64)[:, None]
    return big, cart, sp


def lattice_si(reps=(2, 2, 2), a=5.431):
    cell = np.eye(3) * a
    return _supercell(cell, _diamond_basis(), ["Si"] * 8, reps)


def lattice_tio2(reps=(2, 2, 4), a=4.594, c=2.959, u=0.305):
    cell = np.diag([a, a, c])
    basis = np.array([[0, 0, 0], [.5, .5, .5],
                      [u, u, 0], [1 - u, 1 - u, 0], [.5 + u, .5 - u, .5], [.5 - u, .5 + u, .5]])
    return _supercell(cell, basis, ["Ti", "Ti", "O", "O", "O", "O"], reps)


def lattice_abo4(reps=(3, 3, 1), a=5.243, c=11.376, xyz=(0.1507, 0.0086, 0.2106)):
    """Scheelite-like CaWO4 (I4_1/a, origin choice 2): 4 Ca + 4 W + 16 O per conventional cell."""
    cell = np.diag([a, a, c])
    x, y, z = xyz
    w = [[0, .25, .125], [.5, .25, .375], [.5, .75, .625], [0, .75, .875]]
    ca = [[0, .25, .625], [.5, .25, .875], [.5, .75, .125], [0, .75, .375]]
    gen = [[x, y, z], [-x + .5, -y, z + .5], [-y + .75, x + .25, z + .25], [y + .75, -x + .75, z + .75],
           [-x, -y, -z], [x + .5, y, -z + .5], [y + .25, -x + .75, -z + .75], [-y + .25, x + .25, -z + .25]]
    o = gen + [[p[0] + .5, p[1] + .5, p[2] + .5] for p in gen]
    basis = np.mod(np.array(ca + w + o, dtype=np.float64), 1.0)
    return _supercell(cell, basis, ["Ca"] * 4 + ["W"] * 4 + ["O"] * 16, reps)


# ----------------------------------------------------------------------------------------------
# descriptor rows.  G2: [Rc, fc, eta, Rs]; G4: [Rc, fc, eta, xi, lambda]
def _g2(etas, rss, rc=6.0):
    return [[rc, 0.0, e, r] for e in etas for r in rss]


def _g4(etas, xi_lams, rc=6.0):
    return [[rc, 0.0, e, xi, lam] for e in etas for (xi, lam) in xi_lams]


@dataclasses.dataclass
class Config:
    name: str
    config_id: int
    lattice: tuple            # (cell 3x3, cart Nx3, species list)
    n_structures: int
    G2: List[List[float]]
    G4: List[List[float]]
    hidden: List[int]
    e0: float = -5.0
    rcut: float = 6.0
    sigma: float = 0.08
    G1: List[List[float]] = dataclasses.field(default_factory=list)
    G3: List[List[float]] = dataclasses.field(default_factory=list)
    transfer: str = "tanh"

    @property
    def species(self) -> List[str]:
        return self.lattice[2]

    @property
    def elements(self) -> List[int]:
        return sorted({SYMBOL_TO_Z[s] for s in self.species})

    @property
    def natoms(self) -> int:
        return len(self.species)

    @property
    def NG(self) -> int:
        nt = len(self.elements)
        return nt * (len(self.G1) + len(self.G2)) + nt * (nt + 1) // 2 * (len(self.G3) + len(self.G4))


def config_c2(n_structures=10000) -> Config:
    return Config("C2_si64", 2, lattice_si((2, 2, 2)), n_structures,
                  _g2([0.05, 0.2, 0.8, 3.2, 12.8], [0.0, 2.5]),
                  _g4([0.005, 0.02, 0.08, 0.32, 1.28],
                      [(xi, lam) for xi in (1.0, 2.0, 4.0, 16.0) for lam in (1.0, -1.0)]),
                  [40, 40])


def config_c3(n_structures=20000) -> Config:
    return Config("C3_tio2_96", 3, lattice_tio2((2, 2, 4)), n_structures,
                  _g2([0.05, 0.4, 3.2, 12.8], [0.0, 2.0]),
                  _g4([0.01, 0.1, 0.5], [(xi, lam) for xi in (1.0, 4.0) for lam in (1.0, -1.0)]),
                  [40, 40])


def config_c4(n_structures=50000) -> Config:
    return Config("C4_abo4_216", 4, lattice_abo4((3, 3, 1)), n_structures,
                  _g2([0.05, 0.2, 0.8, 3.2, 12.8], [0.0, 2.5]),
                  _g4([0.01, 0.05, 0.2, 0.6, 1.5], [(1.0, 1.0), (1.0, -1.0), (4.0, 1.0)]),
                  [60, 60, 60])


def config_c5(cells=50) -> Config:
    c2 = config_c2(1)
    return Config("C5_si_md", 5, lattice_si((cells, cells, cells)), 1, c2.G2, c2.G4, c2.hidden)


CONFIGS = {"C2": config_c2, "C3": config_c3, "C4": config_c4, "C5": config_c5}


# ----------------------------------------------------------------------------------------------
def make_structure(cfg: Config, s: int):
    """Return (cell 3x3, cart positions Nx3 wrapped into the cell, Z per atom, target, tag)."""
    cell, ideal, species = cfg.lattice
    n = len(species)
    seed = cfg.config_id * 10 ** 9 + s
    disp = normal(seed, 3 * n, stream=0).reshape(n, 3) * cfg.sigma
    cart = ideal + disp
    inv = np.linalg.inv(cell)
    frac = cart @ inv
    frac -= np.floor(frac)
    cart = frac @ cell
    target = cfg.e0 * n + 0.5 * float(normal(seed, 1, stream=7)[0])
    tag = TAG_VAL if (s % 20 == 19) else TAG_TRAIN
    Z = np.array([SYMBOL_TO_Z[x] for x in species], dtype=np.int32)
    return cell.copy(), cart, Z, target, tag


def make_batch(cfg: Config, first: int, count: int):
    """Flat arrays for ``count`` structures starting at index ``first``.

    Returns dict(cells [n,3,3], cart [n*N,3], Z [n*N], natoms [n], targets [n], tags [n]).
    Vectorised over structures (the per-structure RNG streams are independent counters).
    """
    cell, ideal, species = cfg.lattice
    n = len(species)
    Zs = np.array([SYMBOL_TO_Z[x] for x in species], dtype=np.int32)
    inv = np.linalg.inv(cell)
    cart = np.empty((count, n, 3), dtype=np.float64)
    targets = np.empty(count, dtype=np.float64)
    for k in range(count):
        seed = cfg.config_id * 10 ** 9 + first + k
        disp = normal(seed, 3 * n, stream=0).reshape(n, 3) * cfg.sigma
        c = ideal + disp
        f = c @ inv
        f -= np.floor(f)
        cart[k] = f @ cell
        targets[k] = cfg.e0 * n + 0.5 * float(normal(seed, 1, stream=7)[0])
    idx = np.arange(first, first + count)
    tags = np.where(idx % 20 == 19, TAG_VAL, TAG_TRAIN).astype(np.int32)
    return dict(cells=np.broadcast_to(cell, (count, 3, 3)).copy(), cart=cart.reshape(-1, 3),
                Z=np.tile(Zs, count), natoms=np.full(count, n, dtype=np.int32), targets=targets, tags=tags)


def _fmt(v: float) -> str:
    return repr(float(v))   # shortest round-trip decimal: the reference's operator>> recovers the same double


def write_prophet_xml(path: str, structures) -> None:
    """structures: iterable of (cell, cart, Z, target, tag).  Line-oriented like doc/PROPhet.xml."""
    structures = list(structures)
    with open(path, "w") as f:
        f.write("<PROPhet>\n")
        f.write("  <nsystem>%d</nsystem>\n" % len(structures))
        f.write("  <systems>\n")
        for k, (cell, cart, Z, target, tag) in enumerate(structures):
            f.write('    <system id="%d">\n' % (k + 1))
            f.write("      <train>%s</train>\n" % TAG_NAMES[int(tag)])
            f.write('      <lattice units="angstrom">\n')
            for name, row in zip("abc", cell):
                f.write("        <%s>%s %s %s</%s>\n" % (name, _fmt(row[0]), _fmt(row[1]), _fmt(row[2]), name))
            f.write("      </lattice>\n")
            f.write('      <atoms units="angstrom">\n')
            f.write("        <natoms>%d</natoms>\n" % len(Z))
            f.write("        <species>%d</species>\n" % len(set(int(z) for z in Z)))
            for z, r in zip(Z, cart):
                f.write('        <atom specie="%s">%s %s %s</atom>\n'
                        % (Z_TO_SYMBOL[int(z)], _fmt(r[0]), _fmt(r[1]), _fmt(r[2])))
            f.write("      </atoms>\n")
            f.write("      <target>%s</target>\n" % _fmt(target))
            f.write("    </system>\n")
        f.write("  </systems>\n")
        f.write("</PROPhet>\n")


def write_input_file(path: str, cfg: Config, *, niterations=1, checkpoint_in=None, checkpoint_out="potential",
                     algorithm="rprop", extra: Sequence[str] = (), data: Sequence[str] = ("data code=prophet",)) -> None:
    """An input_file for the reference CLI (keys: Setup.cpp:182-510)."""
    lines = ["network_type = nn", "hidden = " + " ".join(str(h) for h in cfg.hidden),
             "input = structure", "output = energy", "transfer = %s" % cfg.transfer,
             "Rcut = %s" % _fmt(cfg.rcut)]
    for r in cfg.G1:
        lines.append("G1 " + " ".join(_fmt(v) for v in r))
    for r in cfg.G2:
        lines.append("G2 " + " ".join(_fmt(v) for v in r))
    for r in cfg.G3:
        lines.append("G3 " + " ".join(_fmt(v) for v in r))
    for r in cfg.G4:
        lines.append("G4 " + " ".join(_fmt(v) for v in r))
    lines += ["algorithm = %s" % algorithm, "Niterations = %d" % niterations, "threshold = 0", "Nprint = 1",
              "Ncheckpoint = 1000000", "checkpoint_out = %s" % checkpoint_out]
    if checkpoint_in:
        lines.append("checkpoint_in = %s" % checkpoint_in)
    lines += list(extra)
    lines += list(data)     # the data block: "data code=prophet" (systems come from PROPhet.xml) or code=vasp + directories
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")


def write_vasprun(dirpath: str, cell: np.ndarray, frac: np.ndarray, Z: Sequence[int], target: float) -> None:
    """A minimal vasprun.xml holding exactly what the reference's VASP reader looks for (VASP.cpp:78-176):
    atominfo/array[@name=atoms] symbols, structure/crystal basis rows, fractional positions, e_wo_entrp."""
    os.makedirs(dirpath, exist_ok=True)
    with open(os.path.join(dirpath, "vasprun.xml"), "w") as f:
        f.write("<modeling>\n <atominfo>\n  <atoms>  %d </atoms>\n  <types> %d </types>\n  <array name=\"atoms\" >\n   <set>\n"
                % (len(Z), len(set(int(z) for z in Z))))
        for z in Z:
            f.write("    <rc><c>%-2s</c><c>   1</c></rc>\n" % Z_TO_SYMBOL[int(z)])
        f.write("   </set>\n  </array>\n </atominfo>\n <structure>\n  <crystal>\n   <varray name=\"basis\" >\n")
        for row in np.asarray(cell).reshape(3, 3):
            f.write("    <v> %r %r %r </v>\n" % tuple(float(v) for v in row))
        f.write("   </varray>\n  </crystal>\n  <varray name=\"positions\" >\n")
        for row in np.asarray(frac).reshape(-1, 3):
            f.write("   <v> %r %r %r </v>\n" % tuple(float(v) for v in row))
        f.write("  </varray>\n </structure>\n<i name=\"e_wo_entrp\">  %r </i>\n</modeling>\n" % float(target))


def random_network(cfg: Config, Z: int, seed: int) -> Dict[str, np.ndarray]:
    """Random parameters for one element's net, layer-major / node-major / [w..., b] (NN.cpp:83-94).

    Scaled like N(0, 1/sqrt(n_in)) so tanh units stay out of saturation for any descriptor scale."""
    sizes = [cfg.NG] + list(cfg.hidden) + [1]
    parts = []
    stream = 11
    for n_in, n_out in zip(sizes[:-1], sizes[1:]):
        w = normal(seed * 1000 + Z, n_out * (n_in + 1), stream=stream).reshape(n_out, n_in + 1)
        w[:, :n_in] *= 1.0 / np.sqrt(n_in)
        w[:, n_in] *= 0.1
        parts.append(w.reshape(-1))
        stream += 1
    return dict(params=np.concatenate(parts), sizes=np.array(sizes, dtype=np.int32))


# ----------------------------------------------------------------------------------------------
# MD harness helpers: what LAMMPS would hand to pair_style nn (pair_nn.cpp:112-153)
def expand_ghosts(cell: np.ndarray, cart: np.ndarray, Z: np.ndarray, halo: float, with_owner: bool = False):
    """Locals followed by periodic-image ghosts lying within ``halo`` of the (orthorhombic) box.

    with_owner=True also returns, per ghost, the index of the local atom it is an image of and its image shift
    (what LAMMPS' reverse communication uses to return ghost forces to their owners)."""
    L = np.diag(cell)
    assert np.allclose(cell, np.diag(L)), "ghost expansion implemented for orthorhombic boxes"
    nrep = np.ceil(halo / L).astype(int)
    xs, zs, owners, shifts = [cart], [Z], [], []
    idx = np.arange(len(cart))
    for i in range(-nrep[0], nrep[0] + 1):
        for j in range(-nrep[1], nrep[1] + 1):
            for k in range(-nrep[2], nrep[2] + 1):
                if i == 0 and j == 0 and k == 0:
                    continue
                sh = cart + np.array([i, j, k], dtype=np.float64) * L
                keep = np.all((sh > -halo) & (sh < L + halo), axis=1)
                xs.append(sh[keep])
                zs.append(Z[keep])
                owners.append(idx[keep])
                shifts.append(np.tile(np.array([i, j, k], dtype=np.int32), (int(keep.sum()), 1)))
    if with_owner:
        owner = np.concatenate(owners) if owners else np.zeros(0, dtype=np.int64)
        shift = np.concatenate(shifts) if shifts else np.zeros((0, 3), dtype=np.int32)
        return np.concatenate(xs), np.concatenate(zs), len(cart), owner, shift
    return np.concatenate(xs), np.concatenate(zs), len(cart)


def slab_partition(cell: np.ndarray, cart: np.ndarray, rank: int, world: int, halo: float):
    """Rank ``rank``'s share of an orthorhombic periodic box cut into ``world`` slabs along x (what LAMMPS' spatial
    decomposition hands one MPI rank): the atoms it owns, followed by the ghosts -- every periodic image of any atom that
    lies within ``halo`` of the slab (and of the box faces in y and z).

    Returns dict(x [nlocal+nghost,3], nlocal, owned_ids [nlocal] global ids (ascending), ghost_rank [nghost] owning rank
    of each ghost, ghost_local [nghost] its index among the owner's local atoms)."""
    L = np.diag(cell)
    assert np.allclose(cell, np.diag(L)), "slab partition implemented for orthorhombic boxes"
    n = len(cart)
    width = L[0] / world
    owner = np.minimum((cart[:, 0] / width).astype(np.int64), world - 1)
    owner = np.maximum(owner, 0)
    local_of = np.empty(n, dtype=np.int64)
    for r in range(world):
        ids = np.nonzero(owner == r)[0]
        local_of[ids] = np.arange(len(ids))
    owned = np.nonzero(owner == rank)[0]
    xlo, xhi = rank * width, (rank + 1) * width
    nrep = np.ceil(halo / L).astype(int)
    gx, gid = [], []
    idx = np.arange(n)
    for i in range(-nrep[0], nrep[0] + 1):
        for j in range(-nrep[1], nrep[1] + 1):
            for k in range(-nrep[2], nrep[2] + 1):
                sh = cart + np.array([i, j, k], dtype=np.float64) * L
                keep = ((sh[:, 0] > xlo - halo) & (sh[:, 0] < xhi + halo) & (sh[:, 1] > -halo) & (sh[:, 1] < L[1] + halo) &
                        (sh[:, 2] > -halo) & (sh[:, 2] < L[2] + halo))
                if i == 0 and j == 0 and k == 0:
                    keep &= owner != rank
                gx.append(sh[keep])
                gid.append(idx[keep])
    gid = np.concatenate(gid)
    x = np.concatenate([cart[owned]] + gx)
    return dict(x=np.ascontiguousarray(x), nlocal=len(owned), owned_ids=owned, ghost_rank=owner[gid].astype(np.int32),
                ghost_local=local_of[gid].astype(np.int32))


def brute_full_neighbor_list(x: np.ndarray, nlocal: int, cutneigh: float):
    """Full list (locals -> locals+ghosts) with LAMMPS' test rsq < cutneigh^2, ascending j."""
    numneigh = np.zeros(nlocal, dtype=np.int32)
    rows = []
    c2 = cutneigh * cutneigh
    for i in range(nlocal):
        d = x - x[i]
        rsq = d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2]
        j = np.nonzero(rsq < c2)[0]
        j = j[j != i].astype(np.int32)
        rows.append(j)
        numneigh[i] = len(j)
    offsets = np.zeros(nlocal + 1, dtype=np.int64)
    offsets[1:] = np.cumsum(numneigh)
    flat = np.concatenate(rows) if rows else np.zeros(0, dtype=np.int32)
    return numneigh, offsets, flat.astype(np.int32)
