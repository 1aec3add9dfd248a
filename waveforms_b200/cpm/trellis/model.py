"""Trellis descriptions -- host-side mirror of waveforms/cpm/trellis/model.py:10-293.

Same public objects (Branch, Trellis, FiniteStateMachine, the five trellis
constants, forward_map/reverse_map/filter_branches), but the SOQPSK tables are
*generated* from the state convention instead of being listed: state = 2*I + Q
(latest even-indexed bit, latest odd-indexed bit); even columns update I, odd
columns update Q; alpha = 2(-1)^(i+1)(2*other-1)(new-old).  Branch order inside a
column (start ascending, then new bit 0/1) is semantically significant for the
detector's tie-break (viterbi/algorithm.py:77-84) and matches the reference.
``flatten_forward`` turns any Trellis into the int8 tables wf_trellis_encode takes.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass
class Branch:
    inp: int
    out: int
    start: int
    end: int


@dataclass
class Trellis:
    branches: list[list[Branch]]

    @property
    def input_cardinality(self) -> int:
        """Bits per input value; ValueError unless the input alphabet is a power of two."""
        n_inputs = len({b.inp for col in self.branches for b in col})
        bits = n_inputs.bit_length() - 1
        if n_inputs != 1 << bits:
            err = "Number of unique inputs should be multiple of 2."
            raise ValueError(err)
        return bits

    @property
    def output_cardinality(self) -> int:
        return len({b.out for col in self.branches for b in col})

    @property
    def columns(self) -> int:
        return len(self.branches)

    @property
    def branches_per_column(self) -> int:
        return len(self.branches[0])

    @property
    def states(self) -> int:
        return 1 + max(b.start for col in self.branches for b in col)


def filter_branches(state: int, branches: list[Branch], attr: str = "start") -> list[Branch]:
    return [b for b in branches if getattr(b, attr) == state]


def forward_branches(state: int, branches: list[Branch]) -> list[Branch]:
    return filter_branches(state, branches, "start")


def reverse_branches(state: int, branches: list[Branch]) -> list[Branch]:
    return filter_branches(state, branches, "end")


def forward_map(state: int, branches: list[Branch]) -> dict[int, Branch]:
    """input -> branch leaving ``state`` (a later duplicate overrides an earlier one)."""
    return {b.inp: b for b in forward_branches(state, branches)}


def reverse_map(state: int, branches: list[Branch]) -> dict[int, Branch]:
    """output -> branch entering ``state``."""
    return {b.out: b for b in reverse_branches(state, branches)}


class FiniteStateMachine:
    def __init__(self, trellis: Trellis) -> None:
        self.trellis = trellis
        self.branches_per_column = trellis.branches_per_column
        self.columns = trellis.columns
        self.states = trellis.states
        rng = range(self.states)
        self.forward_branch_mapping = [[forward_map(s, col) for s in rng] for col in trellis.branches]
        self.reverse_branch_mapping = [[reverse_map(s, col) for s in rng] for col in trellis.branches]
        self.forward_transitions = [[{b.end: b for b in forward_branches(s, col)} for s in rng] for col in trellis.branches]
        self.reverse_transitions = [[{b.start: b for b in reverse_branches(s, col)} for s in rng] for col in trellis.branches]
        self.symbols = sorted({b.out for col in trellis.branches for b in col})
        self.symbol_idx_map = {sym: i for i, sym in enumerate(self.symbols)}


def _soqpsk_column(parity: int, differential: bool) -> list[Branch]:
    col = []
    for start in range(4):
        i_bit, q_bit = start >> 1, start & 1
        old, other = (i_bit, q_bit) if parity == 0 else (q_bit, i_bit)
        sign = -1 if parity == 0 else 1
        for new in (0, 1):
            end = (new << 1) | q_bit if parity == 0 else (i_bit << 1) | new
            col.append(Branch(inp=(new ^ old) if differential else new,
                              out=2 * sign * (2 * other - 1) * (new - old), start=start, end=end))
    return col


SOQPSKTrellis4x2 = Trellis(branches=[_soqpsk_column(0, False), _soqpsk_column(1, False)])
SOQPSKTrellis4x2DiffEncoded = Trellis(branches=[_soqpsk_column(0, True), _soqpsk_column(1, True)])
# the same machine unrolled into one 8-state column: states 0-3 take the even-column
# branch into 4+end, states 4-7 the odd-column branch back into 0-3
SOQPSKTrellis8x1 = Trellis(branches=[
    [Branch(b.inp, b.out, b.start, 4 + b.end) for b in _soqpsk_column(0, False)]
    + [Branch(b.inp, b.out, 4 + b.start, b.end) for b in _soqpsk_column(1, False)],
])


def _memoryless(m: int) -> Trellis:
    return Trellis(branches=[[Branch(inp=v, out=2 * v - (m - 1), start=s, end=v) for s in range(m) for v in range(m)]])


SimpleTrellis2 = _memoryless(2)  # out = 2*in - 1
SimpleTrellis4 = _memoryless(4)  # out = 2*in - 3


def flatten_forward(trellis: Trellis) -> tuple[np.ndarray, np.ndarray]:
    """(next_state, out_symbol) int8[columns][states][2**bits]; -1 marks a missing branch."""
    n_in = 1 << trellis.input_cardinality
    nxt = np.full((trellis.columns, trellis.states, n_in), -1, dtype=np.int8)
    out = np.zeros((trellis.columns, trellis.states, n_in), dtype=np.int8)
    for c, col in enumerate(trellis.branches):
        for b in col:  # later duplicates override, like the dict in forward_map
            nxt[c, b.start, b.inp] = b.end
            out[c, b.start, b.inp] = b.out
    return nxt, out


def same_trellis(a: Trellis, b: Trellis) -> bool:
    return a.branches == b.branches
