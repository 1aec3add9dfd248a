"""Float64 CPU restatement of the reference hot path (TEST INFRASTRUCTURE ONLY).

Every function cites the reference lines (relative to /root/reference) whose
behaviour it restates.  Written from the behaviour, not copied: data tables are
re-derived or re-encoded, loops are re-expressed.  Nothing under
``waveforms_b200`` may import this module.
"""
from __future__ import annotations

import numpy as np

# --------------------------------------------------------------------------
# A1  PRBS: Galois LFSR (waveforms/glfsr/glfsr.py:6-19, glfsr/pn.py:6-107)
# --------------------------------------------------------------------------
# XAPP052 tap sets, degree -> taps; compact restatement of pn.py:6-72.
_TAPS = {}
for _line in (
    "2:2,1 3:3,2 4:4,3 5:5,3 6:6,5 7:7,6 8:8,6,5,4 9:9,5 10:10,7 11:11,9 "
    "12:12,6,4,1 13:13,4,3,1 14:14,5,3,1 15:15,14 16:16,15,13,4 17:17,14 "
    "18:18,11 19:19,6,2,1 20:20,17 21:21,19 22:22,21 23:23,18 24:24,23,22,17 "
    "25:25,22 26:26,6,2,1 27:27,5,2,1 28:28,25 29:29,27 30:30,6,4,1 31:31,28 "
    "32:32,22,2,1 33:33,20 34:34,27,2,1 35:35,33 36:36,25 37:37,5,4,3,2,1 "
    "38:38,6,5,1 39:39,35 40:40,38,21,19 41:41,38 42:42,41,20,19 "
    "43:43,42,38,37 44:44,43,18,17 45:45,44,42,41 46:46,45,26,25 47:47,42 "
    "48:48,47,21,20 49:49,40 50:50,49,24,23 51:51,50,36,35 52:52,49 "
    "53:53,52,38,37 54:54,53,18,17 55:55,31 56:56,55,35,34 57:57,50 58:58,39 "
    "59:59,58,38,37 60:60,59 61:61,60,46,45 62:62,61,6,5 63:63,62 "
    "64:64,63,61,60"
).split():
    _d, _t = _line.split(":")
    _TAPS[int(_d)] = tuple(int(x) for x in _t.split(","))


def pn_mask(degree: int) -> int:
    """pn.py:75-90 -- mask = sum 1<<(tap-1); KeyError unless 1 < degree < 65."""
    if not 1 < degree < 65:
        raise KeyError(f"PRBS Polynomial Not Defined for {degree}.")
    return sum(1 << (t - 1) for t in _TAPS[degree])


def lfsr_bits(mask: int, state: int, count: int) -> tuple[np.ndarray, int]:
    """glfsr.py:15-19 stepped ``count`` times; returns (bits uint8, new state)."""
    out = np.empty(count, dtype=np.uint8)
    for i in range(count):
        b = state & 1
        state >>= 1
        if b:
            state ^= mask
        out[i] = b
    return out, state


def pn_sequence(degree: int) -> np.ndarray:
    """pn.py:93-107 -- one full period (2^n - 1 bits) from the all-ones state."""
    return lfsr_bits(pn_mask(degree), (1 << degree) - 1, (1 << degree) - 1)[0]


def pn_frame_bits(degree: int, frame: int, nbits: int) -> np.ndarray:
    """Frame ``frame`` = stream bits [frame*nbits, (frame+1)*nbits) of the
    periodic PN stream (SURVEY 8d synthetic-input convention)."""
    period = (1 << degree) - 1
    seq = pn_sequence(degree)
    idx = (frame * nbits + np.arange(nbits)) % period
    return seq[idx]


# --------------------------------------------------------------------------
# A2  Trellis tables (waveforms/cpm/trellis/model.py:179-293)
# A branch is (inp, out, start, end); list order inside a column is
# semantically significant for the detector's tie-break (algorithm.py:77-84).
# --------------------------------------------------------------------------
def _soqpsk_columns(diff: bool):
    """Re-derive the two SOQPSK 4-state columns instead of listing them.

    State = 2*I + Q (latest even-index bit, latest odd-index bit).  Even
    columns update I, odd columns update Q.  Non-differential: new bit = inp.
    Differential (model.py:233-258): new bit = inp XOR old bit of that rail.
    Symbol: alpha = 2*(-1)^(i+1)*(2*other-1)*(new-old)  (notes/cpm/soqpsk.md).
    Branch order: for start in 0..3, listed by *new bit* 0 then 1
    (model.py:205-258 lists end-state ascending for each start).
    """
    cols = []
    for parity in (0, 1):
        col = []
        for start in range(4):
            i_bit, q_bit = start >> 1, start & 1
            old, other = (i_bit, q_bit) if parity == 0 else (q_bit, i_bit)
            for new in (0, 1):
                inp = (new ^ old) if diff else new
                sign = -1 if parity == 0 else 1
                out = 2 * sign * (2 * other - 1) * (new - old)
                end = (new << 1 | q_bit) if parity == 0 else (i_bit << 1 | new)
                col.append((inp, out, start, end))
        cols.append(col)
    return cols


def _memoryless(m: int):
    return [[(v, 2 * v - (m - 1), s, v) for s in range(m) for v in range(m)]]


def _soqpsk_8x1():
    """model.py:179-202 -- same machine unrolled to 8 states in ONE column:
    states 0-3 take the even-column branch to 4+end, states 4-7 the odd one."""
    e, o = _soqpsk_columns(False)
    col = [(i, a, s, 4 + t) for (i, a, s, t) in e]
    # reference 8x1 numbers the mid states by (end-state order of the 4x2
    # trellis) so that 4+end is reached; odd half returns to 0..3.
    col += [(i, a, 4 + s, t) for (i, a, s, t) in o]
    return [col]


TRELLISES = {
    "SOQPSKTrellis4x2": _soqpsk_columns(False),
    "SOQPSKTrellis4x2DiffEncoded": _soqpsk_columns(True),
    "SOQPSKTrellis8x1": _soqpsk_8x1(),
    "SimpleTrellis2": _memoryless(2),
    "SimpleTrellis4": _memoryless(4),
}


def trellis_bits_per_symbol(cols) -> int:
    """model.py:22-36 input_cardinality."""
    n_in = len({b[0] for c in cols for b in c})
    if 1 << (n_in.bit_length() - 1) != n_in:
        raise ValueError("Number of unique inputs should be multiple of 2.")
    return n_in.bit_length() - 1


# --------------------------------------------------------------------------
# A3  TrellisEncoder (waveforms/cpm/trellis/encoder.py:7-59)
# --------------------------------------------------------------------------
def fsm_encode(cols, bits, i0: int = 0, state0: int = 0, return_state=False):
    """Sequential FSM walk: ``card`` bits MSB-first per symbol, column = i %
    columns, follow the branch with matching (start, inp) (encoder.py:33-46).
    NOTE forward_map (model.py:127-137) is a dict keyed by inp, so when two
    branches share (start, inp) the LAST one in list order wins."""
    card = trellis_bits_per_symbol(cols)
    bits = np.asarray(bits)
    if bits.size % card:
        raise ValueError("Input length must be a multiple of FSM cardinality.")
    fwd = []
    for col in cols:
        m = {}
        for (inp, out, s, e) in col:
            m[(s, inp)] = (out, e)
        fwd.append(m)
    nsym = bits.size // card
    out = np.zeros(nsym, dtype=np.int8)
    i, st = i0, state0
    for n in range(nsym):
        v = 0
        for x in range(card):
            v = (v << 1) | int(bits[n * card + x])
        o, st = fwd[i % len(cols)][(st, v)]
        out[n] = o
        i += 1
    if return_state:
        return out, i, st
    return out


# A4 mappers
def map_pcmfm(bits):
    """pcmfm/precoder.py:15 -- 2b-1."""
    return (2 * np.asarray(bits).astype(np.int8) - 1).astype(np.int8)


def map_multih(bits):
    """multih/precoder.py:18-23 (fresh object, even length): 2(2b0+b1)-3."""
    bits = np.asarray(bits)
    if bits.size % 2:
        raise ValueError("Odd length bit array passed into quaternary mapper.")
    return (2 * (2 * bits[0::2] + bits[1::2]).astype(np.int8) - 3).astype(np.int8)


def soqpsk_precoder(bits):
    """soqpsk/precoder.py:19-24 with fresh memory: ternary {-1,0,+1}."""
    a = np.concatenate(([0, 0], np.asarray(bits))).astype(np.int8)
    sgn = np.ones(len(bits), dtype=np.int8)
    sgn[0::2] = -1
    return (sgn * (2 * a[1:-1] - 1) * (a[:-2] - a[2:])).astype(np.int8)


# --------------------------------------------------------------------------
# A5  Frequency pulses (soqpsk/pulse_filters.py:9-116, pcmfm/pulse_filters.py:8-25,
#     multih/pulse_filters.py:11-23, cpm/helpers.py:5-19)
# --------------------------------------------------------------------------
def _half_area(sps: int, g):
    """helpers.py:18-19 -- scale so that sum(g) = sps/2."""
    return (sps / (np.sum(g) * 2)) * g


def pulse_soqpsk(t1=1.5, t2=0.5, rho=0.7, b=1.25, sps=8):
    """soqpsk/pulse_filters.py:29-50: windowed cos*sinc pulse on
    [-2(t1+t2), 2(t1+t2)] sampled at int(4(t1+t2)sps)+1 points."""
    span = 2 * (t1 + t2)
    t = np.linspace(-span, span, num=int(span * sps * 2) + 1, dtype=np.float64)
    x = rho * b * t
    g = np.cos(np.pi * x / 2) / (1 - np.power(x, 2)) * np.sinc(b * t / 2)
    w = np.ones_like(t)
    if t2 > 0:
        edge = (np.abs(t) >= 2 * t1) & (np.abs(t) <= span)
        w[edge] = (1 + np.cos(np.pi * (t[edge] / 2 - t1) / t2)) / 2
        w[np.abs(t) > span] = 0
    gw = g * w
    return (sps / (np.sum(gw) * 2)) * g * w


def pulse_soqpsk_tg(sps=8):
    return pulse_soqpsk(sps=sps)  # :107-116


def pulse_soqpsk_a(sps=8):
    return pulse_soqpsk(b=1.35, t1=1.4, t2=0.6, rho=1.0, sps=sps)  # :53-62


def pulse_soqpsk_b(sps=8):
    return pulse_soqpsk(b=1.45, t1=2.8, t2=1.2, rho=0.5, sps=sps)  # :65-74


def pulse_soqpsk_mil(sps=8):
    g = np.full(sps + 1, 0.5)  # :102-104
    g[0] = 0.0
    return g


def pulse_pcmfm(sps=8, order=4):
    """pcmfm/pulse_filters.py:18-25: NRZ rect(sps)/(2sps) convolved with the
    impulse response of a magnitude-normalised Bessel low-pass."""
    from scipy.signal import besselap, impulse

    span = 3
    tt = np.linspace(0, span * 2 / 0.7, num=(span - 1) * sps + 1, endpoint=False)
    _, h = impulse(besselap(order, norm="mag"), T=tt)
    return _half_area(sps, np.convolve(np.ones(sps) / (2 * sps), h, mode="full"))


def pulse_multih(sps=8, length=3):
    """multih/pulse_filters.py:21-23: raised cosine of ``length`` symbols."""
    t = np.linspace(0, length, num=length * sps + 1)
    return _half_area(sps, (1 - np.cos(2 * np.pi * t / length)) / (2 * length))


# --------------------------------------------------------------------------
# A6  CPM modulator (waveforms/cpm/modulate.py:28-101)
# --------------------------------------------------------------------------
def cpm_modulate(symbols, mod_index, pulse, sps=8, initial_phase=np.pi / 4):
    """modulate.py:81-101 restated step for step (sequential accumulator with
    the per-sample ``% sps`` of :51-53).  Returns (normalized_time, signal)."""
    symbols = np.asarray(symbols)
    h = np.atleast_1d(np.asarray(mod_index, dtype=np.float64))
    n_sym = symbols.size
    n_pts = (n_sym + 1) * sps
    t = np.linspace(0, n_sym + 1, num=n_pts, dtype=np.float64, endpoint=False)
    hk = h[np.arange(n_sym) % h.size]
    up = np.zeros(n_pts, dtype=np.float64)
    up[sps:-1:sps] = symbols * hk
    f = np.convolve(up, pulse, mode="same")
    phase = np.empty(n_pts, dtype=np.float64)
    k = 2 * np.pi / sps
    revs = 0
    for i in range(n_pts):
        revs = (revs + f[i]) % sps
        phase[i] = revs * k + initial_phase
    return t, np.exp(1j * phase)


def cpm_modulate_closed_form(symbols, mod_index, pulse, sps=8, initial_phase=np.pi / 4):
    """Closed form of the above, vectorised; equal to the
    sequential version to ~1e-12.  This is the spec the CUDA modulator follows:
      revs[n] = sum_k a_k h_k (Q[clip(n+off-m_k+1,0,K)] - Q[clip(off-m_k,0,K)])
    with Q the cumulative pulse, off=(K-1)//2, m_k=(k+1)sps."""
    symbols = np.asarray(symbols)
    h = np.atleast_1d(np.asarray(mod_index, dtype=np.float64))
    K = pulse.size
    off = (K - 1) // 2
    n_sym = symbols.size
    n_pts = (n_sym + 1) * sps
    Q = np.concatenate(([0.0], np.cumsum(pulse)))
    w = symbols * h[np.arange(n_sym) % h.size]
    m = (np.arange(n_sym) + 1) * sps
    n = np.arange(n_pts)
    # constant frame-start truncation term (only symbols with m_k < off)
    c0 = float(np.sum(w * Q[np.clip(off - m, 0, K)]))
    # banded evaluation: symbols older than J periods contribute w_k*Q[K]
    J = -(-K // sps) + 1
    k_hi = (n + off) // sps - 1
    full = np.concatenate(([0.0], np.cumsum(w))) * Q[K]
    revs = full[np.clip(k_hi - J + 1, 0, n_sym)] - c0
    for j in range(J):
        k = k_hi - j
        ok = (k >= 0) & (k < n_sym)
        kk = np.clip(k, 0, n_sym - 1)
        idx = np.clip(n + off + 1 - (kk + 1) * sps, 0, K)
        revs = revs + np.where(ok, w[kk] * Q[idx], 0.0)
    return np.exp(1j * (2 * np.pi / sps * revs + initial_phase))


# --------------------------------------------------------------------------
# A7  AWGN (waveforms/noise.py:8-32; Eb/N0 relation soqpsk_detection.py:132)
# --------------------------------------------------------------------------
def complex_awgn(sigma, size, rng):
    """noise.py:23-32: interleaved (re, im) normal draws, sigma per component."""
    return rng.normal(loc=0, scale=sigma, size=(size, 2)).view(np.complex128).flatten()


def sigma_from_ebn0_db(ebn0_db, sps, bits_per_symbol=1):
    """Invert soqpsk_detection.py:132 (ebn0 = 10log10(sps/(2 sigma^2))) and
    extend to k bits/symbol (build's definition)."""
    return np.sqrt(sps / (2.0 * bits_per_symbol * 10.0 ** (np.asarray(ebn0_db) / 10.0)))


# --------------------------------------------------------------------------
# A8  PT bank, A9 PAM bank (examples/soqpsk_detection.py:57-63,134-173;
#     waveforms/cpm/pamapprox.py:32-102)
# --------------------------------------------------------------------------
def pt_taps(pulse, mod_index, sps):
    """soqpsk_detection.py:135-156: 3 complex (sps+1)-tap filters for alpha in
    (-2, 0, +2) from the middle symbol of the phase pulse."""
    L = int(pulse.size / sps)
    q = np.cumsum(pulse) / sps
    seg = q[int((L - 1) * sps / 2): int((L + 1) * sps / 2) + 1]
    return np.stack([np.exp(-2j * np.pi * mod_index * a * seg) for a in (-2, 0, 2)])


def pt_bank(rx, pulse, mod_index, sps):
    taps = pt_taps(pulse, mod_index, sps)
    return np.stack([np.convolve(rx, c, mode="same") for c in taps])


def pam_unit_pulse(q, h):
    """pamapprox.py:45-49 (the ``2`` variant): 2*len-1 samples."""
    pih = h * np.pi
    u = np.zeros(2 * q.size - 1)
    u[: q.size] = np.sin(2 * pih * q) / np.sin(pih)
    u[q.size:] = np.sin(pih - 2 * pih * q[1:]) / np.sin(pih)
    return u


def rho_pulses(pulse, mod_index, sps, k_max=2):
    """pamapprox.py:73-102 (Perrins eq. 3.21) restated: rho_k = C_k *
    prod_c u(t - A[k,c] T) over 2L shifted copies, cropped."""
    L = int(pulse.size / sps)
    u = pam_unit_pulse(np.cumsum(pulse) / sps, mod_index)
    out = []
    for k in range(k_max):
        # row k of the shift matrix: x + c*(k>0), x in 0..L-1, c in {0,1}
        shifts = [x + c * (1 if k > 0 else 0) for x in range(L) for c in range(2)]
        prod = np.ones(u.size + L * sps)
        for s in shifts:
            prod = prod * np.concatenate((np.zeros(s * sps), u, np.zeros((L - s) * sps)))
        out.append(((k + 1.0) * prod)[shifts[-1] * sps: -L * sps])
    return out


PSEUDO_SYMBOLS = np.array(
    [[-1j, 1, 1j], [np.sqrt(2) / 2 * (1 - 1j), np.sqrt(2) / 2, np.sqrt(2) / 2 * (1 + 1j)]],
    dtype=np.complex128,
)  # soqpsk_detection.py:57-63


def pam_bank(rx, pulse, mod_index, sps):
    """soqpsk_detection.py:159-173."""
    rho = rho_pulses(pulse, mod_index, sps, 2)
    d = max(r.size for r in rho)
    out = np.zeros((3, rx.size), dtype=np.complex128)
    for s in range(3):
        for k in range(2):
            rk = np.concatenate((rho[k], np.zeros(d - rho[k].size)))
            out[s] += np.convolve(rx, rk, mode="same") * np.conj(PSEUDO_SYMBOLS[k, s])
    return out


def symbol_instants(n_samples, sps, timing_offset, length=2):
    """soqpsk_detection.py:189-192: n in [0, size - length*sps) with
    (n + timing_offset) % sps == 0."""
    n = np.arange(n_samples - length * sps)
    return n[(n + timing_offset) % sps == 0]


# --------------------------------------------------------------------------
# A10  SOQPSK trellis detector (waveforms/viterbi/algorithm.py:19-101)
# --------------------------------------------------------------------------
_STATE_ROT = (1j, -1.0, 1.0, -1j)  # algorithm.py:30
_ROW = {-2: 0, 0: 1, 2: 2}         # model.py:175-176 (sorted symbols)


class SoqpskDetector:
    """Stateful per-call restatement of algorithm.py:44-101 for any ``length``.

    Carried state: call counter i, the last ``length`` columns of 8 branch
    increments, and the survivor metrics of the OLDEST window stage.  Each
    call redoes ``length`` add-compare-select stages over the window and
    traces back from argmin; strict '<' keeps the first minimum in branch-list
    order (:82); np.argmin keeps the lowest state on ties (:92)."""

    def __init__(self, length=2, diff=True):
        self.cols = TRELLISES["SOQPSKTrellis4x2DiffEncoded" if diff else "SOQPSKTrellis4x2"]
        self.length = length
        self.i = 0
        self.inc = np.zeros((8, length))
        self.carry = np.zeros(4)

    def step(self, z):
        L = self.length
        col_now = self.cols[self.i % 2]
        self.inc[:, :-1] = self.inc[:, 1:]
        self.inc[:, -1] = [np.real(_STATE_ROT[s] * z[_ROW[o]]) for (_, o, s, _e) in col_now]
        prev = self.carry - self.carry.min()
        met = np.zeros((4, L))
        src = np.zeros((4, L), dtype=np.int64)
        for j in range(L):
            col = self.cols[(self.i + j - 1) % 2]
            for st in range(4):
                best, arg = np.inf, 0
                for b, (_i, _o, s, e) in enumerate(col):
                    if e != st:
                        continue
                    m = prev[s] + self.inc[b, j]
                    if m < best:
                        best, arg = m, s
                met[st, j], src[st, j] = best, arg
            prev = met[:, j]
        self.carry = met[:, 0].copy()
        bits = np.zeros(L)
        syms = np.zeros(L)
        st = int(np.argmin(met[:, -1]))
        for j in range(L - 1, -1, -1):
            col = self.cols[(self.i + j - 1) % 2]
            p = src[st, j]
            # reverse_transitions (model.py:170-173) is a dict keyed by start:
            # the LAST listed branch with (start=p, end=st) wins.
            br = [b for b in col if b[2] == p and b[3] == st][-1]
            bits[j], syms[j] = br[0], br[1]
            st = p
        self.i += 1
        return bits, syms


def detect_soqpsk(z_calls, length=2, diff=True):
    """Driver loop soqpsk_detection.py:189-198: feed one MF 3-vector per call,
    keep element [0] of each call's output.  z_calls: complex [calls, 3]."""
    det = SoqpskDetector(length, diff)
    n = z_calls.shape[0]
    bits = np.zeros(n, dtype=np.uint8)
    syms = np.zeros(n, dtype=np.int8)
    for c in range(n):
        b, s = det.step(z_calls[c])
        bits[c], syms[c] = b[0], s[0]
    return bits, syms


def detect_soqpsk_fast(z_calls, diff=True, return_margin=False):
    """length=2 detector in the compact form of SURVEY Appendix A, vectorised
    over a leading batch axis: z_calls complex [..., calls, 3].  Identical
    float64 operation order per frame (Mn = M - min; stage A with previous
    increments; stage B with current; first-min ties) so decisions are
    bit-identical to ``detect_soqpsk`` -- asserted in tests.

    ``return_margin``: also return, per call, the smallest float64 decision margin (the
    |difference| of the two candidates of each of the 8 compare-selects and the gap argmin
    separates) and the largest |candidate| -- the quantities oracle/make_golden.py:acs_margin
    freezes from the reference object (G5/G7/G11 ``*_margin`` / ``*_scale``).  A float32 detector
    may differ from this one only on calls whose margin/scale is at float32 resolution."""
    z = np.asarray(z_calls)
    batch = z.shape[:-2]
    n = z.shape[-2]
    cols = TRELLISES["SOQPSKTrellis4x2DiffEncoded" if diff else "SOQPSKTrellis4x2"]
    # per column, per end state: the two incoming branches in list order
    inc_tab = []
    for col in cols:
        ent = []
        for st in range(4):
            ent.append([(b, s, i, o) for b, (i, o, s, e) in enumerate(col) if e == st])
        inc_tab.append(ent)
    M = np.zeros(batch + (4,))
    bprev = np.zeros(batch + (8,))
    bits = np.zeros(batch + (n,), dtype=np.uint8)
    syms = np.zeros(batch + (n,), dtype=np.int8)
    rot = np.array(_STATE_ROT)
    margin = np.zeros(batch + (n,))
    scale = np.zeros(batch + (n,))
    for c in range(n):
        col = cols[c % 2]
        bcur = np.stack(
            [np.real(rot[s] * z[..., c, _ROW[o]]) for (_i, o, s, _e) in col], axis=-1
        )
        Mn = M - M.min(axis=-1, keepdims=True)
        tabA, tabB = inc_tab[(c - 1) % 2], inc_tab[c % 2]
        M0 = np.empty_like(M)
        p0 = np.zeros(batch + (4,), dtype=np.int64)
        bitA = np.zeros(batch + (4,), dtype=np.uint8)
        symA = np.zeros(batch + (4,), dtype=np.int8)
        for st in range(4):
            (b1, s1, i1, o1), (b2, s2, i2, o2) = tabA[st]
            m1 = Mn[..., s1] + bprev[..., b1]
            m2 = Mn[..., s2] + bprev[..., b2]
            second = m2 < m1
            if return_margin:
                d, a = np.abs(m1 - m2), np.maximum(np.abs(m1), np.abs(m2))
                margin[..., c] = d if st == 0 else np.minimum(margin[..., c], d)
                scale[..., c] = a if st == 0 else np.maximum(scale[..., c], a)
            M0[..., st] = np.where(second, m2, m1)
            p0[..., st] = np.where(second, s2, s1)
            bitA[..., st] = np.where(second, i2, i1)
            symA[..., st] = np.where(second, o2, o1)
        M1 = np.empty_like(M)
        p1 = np.zeros(batch + (4,), dtype=np.int64)
        for st in range(4):
            (b1, s1, _, _), (b2, s2, _, _) = tabB[st]
            m1 = M0[..., s1] + bcur[..., b1]
            m2 = M0[..., s2] + bcur[..., b2]
            second = m2 < m1
            if return_margin:
                margin[..., c] = np.minimum(margin[..., c], np.abs(m1 - m2))
                scale[..., c] = np.maximum(scale[..., c], np.maximum(np.abs(m1), np.abs(m2)))
            M1[..., st] = np.where(second, m2, m1)
            p1[..., st] = np.where(second, s2, s1)
        s1st = np.argmin(M1, axis=-1)
        s0 = np.take_along_axis(p1, s1st[..., None], axis=-1)
        bits[..., c] = np.take_along_axis(bitA, s0, axis=-1)[..., 0]
        syms[..., c] = np.take_along_axis(symA, s0, axis=-1)[..., 0]
        if return_margin:
            fin = np.sort(M1, axis=-1)
            margin[..., c] = np.minimum(margin[..., c], fin[..., 1] - fin[..., 0])
        M, bprev = M0, bcur
    if return_margin:
        return bits, syms, margin, scale
    return bits, syms


# --------------------------------------------------------------------------
# A13  error counting (examples/soqpsk_detection.py:200-209)
# --------------------------------------------------------------------------
def count_errors(dec_bits, dec_syms, tx_bits, tx_syms, length=2):
    """Drop the first ``length`` detector outputs, compare the first
    min(len(tx_syms), remaining) decisions; returns (bit_errs, sym_errs, n)."""
    db = np.asarray(dec_bits)[length:].astype(np.uint8)
    ds = np.asarray(dec_syms)[length:].astype(np.int8)
    n = min(len(tx_syms), len(ds))
    sym_err = int(np.count_nonzero(ds[:n] - np.asarray(tx_syms)[:n]))
    bit_err = int(np.count_nonzero(db[:n] - np.asarray(tx_bits, dtype=np.uint8)[:n]))
    return bit_err, sym_err, n


# --------------------------------------------------------------------------
# End-to-end frame (examples/soqpsk_detection.py:45-216, headless)
# --------------------------------------------------------------------------
def soqpsk_link_frame(bits, pulse, sps, sigma, rng=None, noise=None, detector="PAM",
                      timing_offset=0, mod_index=0.25, length=2, diff=True,
                      fast=True, return_all=False):
    """One frame through the reference chain.  ``noise`` (complex128[T]) may be
    supplied instead of ``rng`` (used by parity tests: frozen PCG64 draws).
    Order follows the script: modulate (:78), draw noise (:84), de-rotate by
    -pi/4 (:85), add (:89), bank (:134-173), detect (:177-198), count (:200-209)."""
    bits = np.asarray(bits, dtype=np.uint8)
    syms = fsm_encode(TRELLISES["SOQPSKTrellis4x2DiffEncoded" if diff else "SOQPSKTrellis4x2"], bits)
    s = cpm_modulate_closed_form(syms, mod_index, pulse, sps) if fast else \
        cpm_modulate(syms, mod_index, pulse, sps)[1]
    if noise is None:
        noise = complex_awgn(sigma, s.size, rng)
    rx = s * np.exp(-1j * np.pi / 4) + noise
    mf = pt_bank(rx, pulse, mod_index, sps) if detector == "PT" else pam_bank(rx, pulse, mod_index, sps)
    inst = symbol_instants(rx.size, sps, timing_offset, length)
    z = mf[:, inst].T
    margin = scale = None
    if fast and length == 2:
        db, ds, margin, scale = detect_soqpsk_fast(z, diff, return_margin=True)
    else:
        db, ds = detect_soqpsk(z, length, diff)
    be, se, n = count_errors(db, ds, bits, syms, length)
    if return_all:
        return dict(bits=bits, symbols=syms, signal=s, rx=rx, z=z, dec_bits=db,
                    dec_syms=ds, bit_errors=be, sym_errors=se, compared=n, margin=margin, scale=scale)
    return be, se, n
