"""Reader / writer for PROPhet's per-element checkpoint files ``<base>_<Symbol>`` (harness side).

Format: Functional_params::print (Functional_params.cpp:133-193) followed by
Neural_network::print (NN.cpp:828-847) / Neural_network_node::print (NN_node.cpp:232-243);
layout documented in SURVEY.md appendix B.  The C++ host library has its own reader/writer for the
product path; this module exists so tests and the golden generator can fabricate checkpoints for the
reference and compare files byte for byte.
"""
from __future__ import annotations

from typing import Dict, List, Sequence

import numpy as np


def _g(v: float, prec: int) -> str:
    """What ``ostream << setprecision(prec) << double`` prints (%g without trailing zeros)."""
    s = "%.*g" % (prec, float(v))
    return s


def write_checkpoint(path: str, symbol: str, all_symbols: Sequence[str], rows: Dict[str, List[List[float]]],
                     means: np.ndarray, scales: np.ndarray, out_mean: float, out_scale: float,
                     sizes: Sequence[int], params: np.ndarray, transfer: str = "tanh",
                     node_transfer: Sequence[str] | None = None, header_precision: int = 12,
                     weight_precision: int = 18, free_energy: float | None = None) -> None:
    """sizes = [NG, hidden..., 1]; params layer-major, node-major, [w_0..w_{n-1}, b] per node."""
    sizes = [int(x) for x in sizes]
    hidden = sizes[1:-1]
    out = []
    out.append("nn")
    out.append("structure  ")
    out.append("%s:  %s " % (symbol, " ".join(all_symbols)))
    if free_energy is not None:      # Functional_params.cpp:148-152
        out.append("FE:  " + _g(free_energy, 9))
    n_rows = sum(len(rows.get(k, [])) for k in ("G1", "G2", "G3", "G4"))
    out.append(str(n_rows))
    for kind in ("G1", "G2", "G3", "G4"):
        for r in rows.get(kind, []):
            out.append(kind + " " + " ".join(_g(v, header_precision) for v in r))
    out.append("".join(_g(v, header_precision) + "  " for v in means))
    out.append("".join(_g(v, header_precision) + "  " for v in scales))
    out.append("energy")
    out.append(_g(out_mean, header_precision))
    out.append(_g(out_scale, header_precision))
    out.append("".join("%d  " % h for h in hidden) if hidden else "0")
    p = 0
    node_index = 0
    n_layers = len(sizes) - 1
    for layer in range(n_layers):
        out.append("[[ layer %d ]] " % layer)
        n_in, n_out = sizes[layer], sizes[layer + 1]
        for _ in range(n_out):
            if layer == n_layers - 1:
                tf = "linear"
            elif node_transfer is not None:
                tf = node_transfer[node_index]
            else:
                tf = transfer
            out.append("  [ node %d ]  %s" % (node_index, tf))
            w = params[p:p + n_in]
            b = params[p + n_in]
            out.append("".join("   " + _g(v, weight_precision) + "  " for v in w))
            out.append("   " + _g(b, weight_precision))
            p += n_in + 1
            node_index += 1
    assert p == len(params), (p, len(params))
    with open(path, "w") as f:
        f.write("\n".join(out) + "\n")


def read_checkpoint(path: str) -> dict:
    """Parse one checkpoint file (mirrors Functional_params::read :207-372 and Neural_network::load :158-200)."""
    with open(path) as f:
        lines = f.read().split("\n")
    it = iter(lines)
    res: dict = {}
    res["network_type"] = next(it).strip()
    res["inputs"] = next(it).split()
    tok = next(it).split()
    res["symbol"] = tok[0].rstrip(":")
    res["all_symbols"] = tok[1:]
    line = next(it)
    if "FE" in line:
        res["FE"] = float(line.split()[1])
        line = next(it)
    n_rows = int(line.split()[0])
    rows = {"G1": [], "G2": [], "G3": [], "G4": []}
    for _ in range(n_rows):
        tok = next(it).split()
        rows[tok[0]].append([float(x) for x in tok[1:]])
    res["rows"] = rows
    res["means"] = np.array([float(x) for x in next(it).split()])
    res["scales"] = np.array([float(x) for x in next(it).split()])
    res["output"] = next(it).strip()
    res["out_mean"] = float(next(it).split()[0])
    res["out_scale"] = float(next(it).split()[0])
    hidden = [int(x) for x in next(it).split()]
    if hidden == [0]:
        hidden = []
    nt = len(res["all_symbols"])
    NG = nt * (len(rows["G1"]) + len(rows["G2"])) + nt * (nt + 1) // 2 * (len(rows["G3"]) + len(rows["G4"]))
    sizes = [NG] + hidden + [1]
    res["sizes"] = sizes
    params: List[float] = []
    transfers: List[str] = []
    rest = list(it)
    k = 0
    for layer in range(len(sizes) - 1):
        while "[[ layer " not in rest[k]:
            k += 1
        k += 1
        for _ in range(sizes[layer + 1]):
            while "[ node " not in rest[k]:
                k += 1
            transfers.append(rest[k].split()[4])
            w = [float(x) for x in rest[k + 1].split()]
            b = float(rest[k + 2].split()[0])
            if len(w) != sizes[layer]:
                raise ValueError("node with %d weights, expected %d" % (len(w), sizes[layer]))
            params.extend(w)
            params.append(b)
            k += 3
    res["params"] = np.array(params)
    res["node_transfer"] = transfers
    return res
