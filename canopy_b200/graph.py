"""Flat gate-DAG description shared by the engine bindings, the generators and the test oracle.

`Graph` is the numpy mirror of `canopy_graph_desc` (include/canopy.h): what a walk over the
reference object model -- Model -> FaultTree::top_events() -> Gate::formula().args()
(reference src/mef/openpsa/fault_tree.h:142, event/gate.h:45-48, event/formula.h:100) -- flattens to.
Connective numbering equals `enum Connective` (reference src/mef/openpsa/event/event.h:143-157).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Iterable, Sequence

import numpy as np

AND, OR, ATLEAST, XOR, NOT, NAND, NOR, NULL, IFF, IMPLY, CARDINALITY = range(11)
CONNECTIVE_NAMES = ("and", "or", "atleast", "xor", "not", "nand", "nor", "null", "iff", "imply", "cardinality")
COMPLEMENT = 0x80000000
ID_MASK = 0x7FFFFFFF
CONST_FALSE = 0x7FFFFFFE
CONST_TRUE = 0x7FFFFFFF


class CGraphDesc(C.Structure):
    _fields_ = [
        ("n_basic", C.c_uint32), ("n_gates", C.c_uint32), ("n_roots", C.c_uint32), ("reserved0", C.c_uint32),
        ("n_edges", C.c_uint64),
        ("p", C.POINTER(C.c_double)), ("op", C.POINTER(C.c_uint8)),
        ("kmin", C.POINTER(C.c_uint16)), ("kmax", C.POINTER(C.c_uint16)),
        ("child_off", C.POINTER(C.c_uint64)), ("child", C.POINTER(C.c_uint32)),
        ("roots", C.POINTER(C.c_uint32)),
    ]


def _ptr(a: np.ndarray, ctype):
    return a.ctypes.data_as(C.POINTER(ctype))


@dataclass
class Graph:
    p: np.ndarray            # float64 [n_basic]
    op: np.ndarray           # uint8   [n_gates]
    kmin: np.ndarray         # uint16  [n_gates]
    kmax: np.ndarray         # uint16  [n_gates]
    child_off: np.ndarray    # uint64  [n_gates+1]
    child: np.ndarray        # uint32  [n_edges]
    roots: np.ndarray        # uint32  [n_roots]
    basic_names: list = field(default_factory=list)
    gate_names: list = field(default_factory=list)
    meta: dict = field(default_factory=dict)

    def __post_init__(self):
        self.p = np.ascontiguousarray(self.p, dtype=np.float64)
        self.op = np.ascontiguousarray(self.op, dtype=np.uint8)
        self.kmin = np.ascontiguousarray(self.kmin, dtype=np.uint16)
        self.kmax = np.ascontiguousarray(self.kmax, dtype=np.uint16)
        self.child_off = np.ascontiguousarray(self.child_off, dtype=np.uint64)
        self.child = np.ascontiguousarray(self.child, dtype=np.uint32)
        self.roots = np.ascontiguousarray(self.roots, dtype=np.uint32)

    @property
    def n_basic(self) -> int:
        return int(self.p.shape[0])

    @property
    def n_gates(self) -> int:
        return int(self.op.shape[0])

    @property
    def n_edges(self) -> int:
        return int(self.child.shape[0])

    @property
    def n_roots(self) -> int:
        return int(self.roots.shape[0])

    def n_tallies(self, tally_all_gates: bool) -> int:
        return self.n_roots + (self.n_gates if tally_all_gates else 0)

    def c_desc(self) -> CGraphDesc:
        """ctypes view; keeps `self` arrays alive only as long as the Graph lives."""
        d = CGraphDesc()
        d.n_basic, d.n_gates, d.n_roots, d.n_edges = self.n_basic, self.n_gates, self.n_roots, self.n_edges
        d.p = _ptr(self.p, C.c_double)
        d.op = _ptr(self.op, C.c_uint8)
        d.kmin = _ptr(self.kmin, C.c_uint16)
        d.kmax = _ptr(self.kmax, C.c_uint16)
        d.child_off = _ptr(self.child_off, C.c_uint64)
        d.child = _ptr(self.child, C.c_uint32)
        d.roots = _ptr(self.roots, C.c_uint32)
        return d

    def nbytes(self) -> int:
        return sum(a.nbytes for a in (self.p, self.op, self.kmin, self.kmax, self.child_off, self.child, self.roots))

    def fanin(self) -> np.ndarray:
        return np.diff(self.child_off.astype(np.int64))


class GraphBuilder:
    """Incremental builder with the reference's vocabulary: basic events, house events (constants),
    gates with a Formula(connective, args, min, max) (reference event/formula.h:72-80)."""

    def __init__(self):
        self._p: list[float] = []
        self._bnames: list[str] = []
        self._gates: list[tuple[int, list[tuple[str, int, bool]], int, int]] = []
        self._gnames: list[str] = []
        self._roots: list[tuple[str, int]] = []

    # handles are ("b", i) / ("g", i) / ("c", 0|1)
    def basic_event(self, p: float, name: str | None = None):
        self._p.append(float(p))
        self._bnames.append(name or f"e{len(self._p) - 1}")
        return ("b", len(self._p) - 1)

    @staticmethod
    def constant(state: bool):
        return ("c", 1 if state else 0)

    def gate(self, connective: int, args: Sequence, min_number: int = 0, max_number: int = 0, name: str | None = None):
        """args: handles or (handle, complement) pairs."""
        norm = []
        for a in args:
            if len(a) == 2 and isinstance(a[0], tuple):
                h, comp = a
            else:
                h, comp = a, False
            norm.append((h[0], h[1], bool(comp)))
        self._gates.append((int(connective), norm, int(min_number), int(max_number)))
        self._gnames.append(name or f"g{len(self._gates) - 1}")
        return ("g", len(self._gates) - 1)

    def set_gate(self, handle, connective: int, args: Sequence, min_number: int = 0, max_number: int = 0):
        """(Re)define a previously created gate: allows forward references and deliberately cyclic inputs."""
        tmp = GraphBuilder()
        tmp.gate(connective, args, min_number, max_number)
        self._gates[handle[1]] = tmp._gates[0]

    def root(self, handle):
        self._roots.append((handle[0], handle[1]))

    def build(self) -> Graph:
        nb = len(self._p)
        op, kmin, kmax, off, child = [], [], [], [0], []
        for conn, args, lo, hi in self._gates:
            op.append(conn); kmin.append(lo); kmax.append(hi)
            for kind, idx, comp in args:
                if kind == "b":
                    cid = idx
                elif kind == "g":
                    cid = nb + idx
                else:
                    cid = CONST_TRUE if idx else CONST_FALSE
                child.append(cid | (COMPLEMENT if comp else 0))
            off.append(len(child))
        roots = [idx if kind == "b" else nb + idx for kind, idx in self._roots]
        return Graph(p=np.array(self._p, dtype=np.float64), op=np.array(op, dtype=np.uint8),
                     kmin=np.array(kmin, dtype=np.uint16), kmax=np.array(kmax, dtype=np.uint16),
                     child_off=np.array(off, dtype=np.uint64), child=np.array(child, dtype=np.uint32),
                     roots=np.array(roots, dtype=np.uint32),
                     basic_names=list(self._bnames), gate_names=list(self._gnames))


def from_lists(p: Iterable[float], gates: Sequence[tuple], roots: Sequence[int]) -> Graph:
    """gates: (op, [child codes], kmin, kmax) with raw uint32 child codes."""
    op, kmin, kmax, off, child = [], [], [], [0], []
    for g in gates:
        o, ch = g[0], g[1]
        lo = g[2] if len(g) > 2 else 0
        hi = g[3] if len(g) > 3 else 0
        op.append(o); kmin.append(lo); kmax.append(hi)
        child.extend(int(c) & 0xFFFFFFFF for c in ch)
        off.append(len(child))
    return Graph(p=np.array(list(p), dtype=np.float64), op=np.array(op, dtype=np.uint8),
                 kmin=np.array(kmin, dtype=np.uint16), kmax=np.array(kmax, dtype=np.uint16),
                 child_off=np.array(off, dtype=np.uint64), child=np.array(child, dtype=np.uint32),
                 roots=np.array(list(roots), dtype=np.uint32))
