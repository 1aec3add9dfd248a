"""DEVELOPMENT TOOL ONLY: run the kernel sources on host threads (build_emu.py) and
compare with the oracle, before spending GPU minutes.  Not a test of the product
library and never imported by it.

    python tools/hostemu/build_emu.py && python tools/hostemu/preflight.py [names...]
"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import cpm_link as o  # noqa: E402
from waveforms_b200 import _ffi  # noqa: E402

L = _ffi.bind(C.CDLL(os.environ.get("WFEMU_LIB", "/tmp/wfemu/libwfb200_emu.so")))
G = lambda n: np.load(os.path.join(ROOT, "tests", "golden", n))  # noqa: E731


def ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def ok(rc):
    assert rc == 0, (rc, L.wf_last_error_string())


def darr(v):
    v = np.ascontiguousarray(np.atleast_1d(v), dtype=np.float64)
    return v, v.ctypes.data_as(C.c_void_p)


def t_lfsr():
    for deg in (2, 9, 15, 23, 33, 64):
        mask = o.pn_mask(deg)
        assert L.wf_lfsr_mask(deg) == mask
        st0 = (1 << deg) - 1
        for k in (0, 1, 31, 1000, 12345):
            _, st = o.lfsr_bits(mask, st0, k)
            assert L.wf_lfsr_jump(mask, deg, st0, k) == st, (deg, k)
    assert L.wf_lfsr_mask(1) == 0 and L.wf_lfsr_mask(65) == 0


def run_prbs(deg, B, N, enc, first=0, state=None, period=True):
    mask, st0 = o.pn_mask(deg), (1 << deg) - 1
    S = N // 2 if enc == 3 else N
    ldw = (N + 31) // 32
    bits = np.full((B, N), 7, np.uint8)
    packed = np.zeros((B, ldw), np.uint32)
    syms = np.full((B, S), 99, np.int8)
    ok(L.wf_prbs_encode(mask, deg, st0, (1 << deg) - 1 if period else 0, None, first, B, N, enc,
                        ptr(state), ptr(bits), ptr(packed), ldw, ptr(syms), None))
    return bits, packed, syms


ENC = {0: "SOQPSKTrellis4x2DiffEncoded", 1: "SOQPSKTrellis4x2", 2: "SimpleTrellis2", 3: "SimpleTrellis4"}


def t_prbs():
    for deg, B, N, first in ((15, 5, 4096, 0), (15, 3, 1000, 7), (9, 2, 70, 3), (23, 2, 2080, 1), (33, 2, 96, 5)):
        for enc in (0, 1, 2, 3, 4):
            bits, packed, syms = run_prbs(deg, B, N, enc, first)
            period = (1 << deg) - 1
            for b in range(B):
                if deg <= 15:
                    ref = o.pn_frame_bits(deg, first + b, N)
                else:
                    st = L.wf_lfsr_jump(o.pn_mask(deg), deg, period, ((first + b) * N) % period)
                    ref, _ = o.lfsr_bits(o.pn_mask(deg), st, N)
                assert np.array_equal(bits[b], ref), (deg, N, enc, b)
                pk = np.packbits(ref, bitorder="little")
                pk = np.concatenate((pk, np.zeros((-pk.size) % 4, np.uint8))).view(np.uint32)
                assert np.array_equal(packed[b], pk), (deg, N, enc, b)
                if enc == 4:
                    rs = o.soqpsk_precoder(ref)
                else:
                    rs = o.fsm_encode(o.TRELLISES[ENC[enc]], ref)
                assert np.array_equal(syms[b], rs), (deg, N, enc, b, np.flatnonzero(syms[b] != rs)[:10])
    # streaming encoder state: two calls == one call
    g = G("g2_encode.npz")
    stream = np.unpackbits(G("g1_pn.npz")["pn15"])
    for n1, n2 in ((1001, 1999), (64, 33), (1, 1)):
        st = np.zeros((1, 2), np.int32)
        a = np.zeros((1, n1), np.int8)
        b = np.zeros((1, n2), np.int8)
        bi = np.ascontiguousarray(stream[:n1][None])
        ok(L.wf_bits_encode(ptr(bi), 1, n1, 0, ptr(st), None, 0, ptr(a), None))
        bi2 = np.ascontiguousarray(stream[n1:n1 + n2][None])
        ok(L.wf_bits_encode(ptr(bi2), 1, n2, 0, ptr(st), None, 0, ptr(b), None))
        ref = g["SOQPSKTrellis4x2DiffEncoded"][:n1 + n2]
        got = np.concatenate((a[0], b[0]))
        assert np.array_equal(got, ref), (n1, n2, np.flatnonzero(got != ref)[:10])
        assert st[0, 0] == n1 + n2
    # generic table-driven trellis
    for name, cols in o.TRELLISES.items():
        card = o.trellis_bits_per_symbol(cols)
        ns = max(max(b[2], b[3]) for c in cols for b in c) + 1
        nxt = np.full((len(cols), ns, 1 << card), -1, np.int8)
        out = np.zeros((len(cols), ns, 1 << card), np.int8)
        for ci, col in enumerate(cols):
            for (inp, ou, s, e) in col:
                nxt[ci, s, inp], out[ci, s, inp] = e, ou
        bi = np.ascontiguousarray(stream[:2048][None])
        sy = np.zeros((1, 2048 // card), np.int8)
        ok(L.wf_trellis_encode(ptr(bi), 1, 2048, ptr(nxt), ptr(out), len(cols), ns, card, None, ptr(sy), None))
        assert np.array_equal(sy[0], g[name][:2048 // card]), name


_MOD = {
    "tg": ("SOQPSKTrellis4x2DiffEncoded", [0.25], o.pulse_soqpsk_tg),
    "pcmfm": ("SimpleTrellis2", [0.7], lambda s: o.pulse_pcmfm(s, 4)),
    "multih": ("SimpleTrellis4", [4 / 16, 5 / 16], o.pulse_multih),
    "mil": ("SOQPSKTrellis4x2DiffEncoded", [0.25], o.pulse_soqpsk_mil),
    "b": ("SOQPSKTrellis4x2", [0.25], o.pulse_soqpsk_b),
}


def modulate(syms, h, pulse, sps, dtype, derot=0.0, sigma=None, seed=0, first=0):
    B, S = syms.shape
    T = (S + 1) * sps
    ld = T + (T & 1)
    out = np.zeros((B, ld), np.complex64 if dtype == 0 else np.complex128)
    hv, hp = darr(h)
    pv, pp = darr(pulse)
    ok(L.wf_cpm_modulate(ptr(syms), B, S, hp, hv.size, pp, pv.size, sps, np.pi / 4, derot, dtype, ptr(out), ld,
                         ptr(sigma), None, seed, first, None, None))
    return out[:, :T]


def t_mod():
    g = G("g4_modulated.npz")
    for lab, (tre, h, pf) in _MOD.items():
        for sps in (8, 10):
            sy = o.fsm_encode(o.TRELLISES[tre], g["bits"])
            ref = g[f"{lab}_{sps}"]
            for dtype, tol in ((0, 1e-5), (1, 1e-10)):
                s = modulate(sy[None], h, pf(sps), sps, dtype)[0]
                err = np.max(np.abs(s - ref))
                print(f"  mod {lab} sps={sps} dtype={dtype} max err {err:.2e}")
                assert err < tol, (lab, sps, dtype, err)
    # several chunks + batch
    bits = np.unpackbits(G("g1_pn.npz")["pn15"])[:2 * 2600].reshape(2, 2600)
    sy = np.stack([o.fsm_encode(o.TRELLISES["SOQPSKTrellis4x2DiffEncoded"], b) for b in bits])
    for dtype, tol in ((0, 1e-5), (1, 1e-10)):
        s = modulate(sy, [0.25], o.pulse_soqpsk_tg(8), 8, dtype)
        for b in range(2):
            ref = o.cpm_modulate_closed_form(sy[b], 0.25, o.pulse_soqpsk_tg(8), 8)
            err = np.max(np.abs(s[b] - ref))
            print(f"  mod long frame {b} dtype={dtype} max err {err:.2e}")
            assert err < tol


def t_modedge():
    """sps-8 fast modulator at degenerate and boundary frame lengths (segment = 5120 intervals)"""
    rng = np.random.default_rng(21)
    for lab, h, pulse, alpha in (("tg", [0.25], o.pulse_soqpsk_tg(8), (-2, 0, 2)),
                                 ("multih", [4 / 16, 5 / 16], o.pulse_multih(8), (-3, -1, 1, 3)),
                                 ("mil", [0.25], o.pulse_soqpsk_mil(8), (-2, 0, 2)),
                                 ("pcmfm", [0.7], o.pulse_pcmfm(8, 4), (-1, 1))):
        for S in (1, 2, 3, 7, 8, 9, 19, 20, 21, 1023, 1024, 1025, 5099, 5115, 5116, 5119, 5120, 5121, 5140, 10241):
            B = 3 if S < 2000 else 1
            sy = rng.choice(np.array(alpha, np.int8), size=(B, S)).astype(np.int8)
            got = modulate(sy, h, pulse, 8, 0)
            sg = np.full(B, 0.8, np.float32)
            buf = np.ascontiguousarray(got.copy())
            ok(L.wf_awgn_add(ptr(buf), 0, B, got.shape[1], got.shape[1], ptr(sg), None, 5, 9, None, None))
            fused = modulate(sy, h, pulse, 8, 0, sigma=sg, seed=5, first=9)
            for b in range(B):
                ref = o.cpm_modulate_closed_form(sy[b], h if len(h) > 1 else h[0], pulse, 8)
                err = np.max(np.abs(got[b] - ref))
                assert err < 1e-5, (lab, S, b, err)
            assert np.max(np.abs(buf - fused)) < 2e-6, (lab, S)
        print(f"  {lab}: all lengths ok")


def mf(rx, taps, comb, sps, n_calls, dtype, first=0):
    B, T = rx.shape
    ld = T
    n_filt, n_taps = taps.shape
    n_out = comb.shape[0] if comb is not None else n_filt
    ldo = n_calls + (n_calls & 1)
    out = np.zeros((B, ldo, n_out), np.complex64 if dtype == 0 else np.complex128)
    rx = np.ascontiguousarray(rx.astype(out.dtype))
    tr, trp = darr(taps.real.ravel())
    ti, tip = darr(taps.imag.ravel())
    if not np.iscomplexobj(taps):
        tip = None
    cr = ci = crp = cip = None
    if comb is not None:
        cr, crp = darr(comb.real.ravel())
        ci, cip = darr(comb.imag.ravel())
    ok(L.wf_mf_bank(ptr(rx), dtype, B, T, ld, trp, tip, n_filt, n_taps, (n_taps - 1) // 2, crp, cip, n_out, sps, first,
                    n_calls, ptr(out), ldo, None))
    return out[:, :n_calls]


def mf_center(rx, taps, sps, n_calls, dtype, first, center):
    B, T = rx.shape
    n_filt, n_taps = taps.shape
    out = np.zeros((B, n_calls, n_filt), np.complex64 if dtype == 0 else np.complex128)
    rxx = np.ascontiguousarray(rx.astype(out.dtype))
    tr, trp = darr(taps.real.ravel())
    ti, tip = darr(taps.imag.ravel())
    ok(L.wf_mf_bank(ptr(rxx), dtype, B, T, T, trp, tip, n_filt, n_taps, center, None, None, n_filt, sps, first, n_calls,
                    ptr(out), n_calls, None))
    return out


def pam_tables(pulse, h, sps):
    rho = o.rho_pulses(pulse, h, sps, 2)
    d = max(r.size for r in rho)
    taps = np.stack([np.concatenate((r, np.zeros(d - r.size))) for r in rho])
    comb = np.conj(o.PSEUDO_SYMBOLS).T  # [3][2]
    return taps, comb


def t_mf():
    g = G("g5_config1.npz")
    pulse = o.pulse_soqpsk_tg(8)
    rx = g["tx"] * np.exp(-1j * np.pi / 4) + g["noise"]
    n_calls = 4095
    for dtype, tol in ((1, 1e-9), (0, 2e-4)):
        z = mf(rx[None], *pam_tables(pulse, 0.25, 8), 8, n_calls, dtype)[0]
        err = np.max(np.abs(z - g["PAM_z"]))
        print(f"  PAM bank dtype={dtype} max err {err:.2e}")
        assert err < tol
        z = mf(rx[None], o.pt_taps(pulse, 0.25, 8), None, 8, n_calls, dtype)[0]
        err = np.max(np.abs(z - g["PT_z"]))
        print(f"  PT bank dtype={dtype} max err {err:.2e}")
        assert err < tol
    # generic shape (sps 8, 5 complex filters of 13 taps, first instant 3)
    rng = np.random.default_rng(0)
    taps = rng.normal(size=(5, 13)) + 1j * rng.normal(size=(5, 13))
    rr = rng.normal(size=(2, 1000)) + 1j * rng.normal(size=(2, 1000))
    z = mf(rr, taps, None, 8, 100, 1, first=3)
    for b in range(2):
        full = np.stack([np.convolve(rr[b], t, "same") for t in taps])
        ref = full[:, 3 + 8 * np.arange(100)].T
        assert np.max(np.abs(z[b] - ref)) < 1e-9


def viterbi(z, dtype, diff=1, length=2, tx_bits=None, tx_syms=None):
    B, n_calls, _ = z.shape
    zz = np.ascontiguousarray(z.astype(np.complex64 if dtype == 0 else np.complex128))
    db = np.zeros((B, n_calls), np.uint8)
    ds = np.zeros((B, n_calls), np.int8)
    cnt = np.zeros((1, 3), np.int64)
    packed = None
    ldw = 0
    S = 0
    if tx_bits is not None:
        S = tx_bits.shape[1]
        pk = np.packbits(tx_bits, axis=1, bitorder="little")
        pk = np.concatenate((pk, np.zeros((B, (-pk.shape[1]) % 4), np.uint8)), axis=1)
        packed = np.ascontiguousarray(pk).view(np.uint32)
        ldw = packed.shape[1]
    ok(L.wf_viterbi_soqpsk(ptr(zz), dtype, B, n_calls, n_calls, diff, length, ptr(db), ptr(ds), ptr(packed), ldw,
                           ptr(tx_syms), S, None, ptr(cnt), None, None))
    return db, ds, cnt


def t_vit():
    g = G("g5_config1.npz")
    for det in ("PT", "PAM"):
        z = g[f"{det}_z"]
        for dtype in (1, 0):
            db, ds, cnt = viterbi(z[None], dtype, tx_bits=g["bits"][None], tx_syms=np.ascontiguousarray(g["symbols"][None]))
            nb = int(np.sum(db[0] != g[f"{det}_dec_bits"]))
            print(f"  viterbi {det} dtype={dtype}: {nb} decision mismatches; counters {cnt[0]}")
            if dtype == 1:
                assert nb == 0
                assert np.array_equal(ds[0], g[f"{det}_dec_syms"])
                assert list(cnt[0]) == [int(g[f"{det}_bit_errors"]), int(g[f"{det}_sym_errors"]), int(g[f"{det}_compared"])]
            else:
                assert nb <= 2
    # other window lengths / non-differential from the G7 variants: regenerate z with the oracle
    g7 = G("g7_variants.npz")
    pulse = o.pulse_soqpsk_tg(8)
    sig = float(np.sqrt(8 / (2 * 10 ** 0.8)))
    for diff in (True, False):
        for length in (2, 4):
            tag = f"d{int(diff)}_l{length}"
            r = o.soqpsk_link_frame(g7["bits"], pulse, 8, sig, noise=g7[f"noise_{tag}"], detector="PAM", length=length,
                                    diff=diff, return_all=True)
            db, ds, cnt = viterbi(r["z"][None], 1, int(diff), length, tx_bits=g7["bits"][None],
                                  tx_syms=np.ascontiguousarray(r["symbols"][None]))
            assert np.array_equal(db[0], g7[f"PAM_bits_{tag}"]), tag
            assert np.array_equal(ds[0], g7[f"PAM_syms_{tag}"]), tag
            assert list(cnt[0]) == list(g7[f"PAM_counts_{tag}"]), (tag, cnt)
    # odd / long windows against the stateful oracle
    z = g["PAM_z"][:300]
    for length in (1, 3, 6, 16):
        rb, rs = o.detect_soqpsk(z, length, True)
        db, ds, _ = viterbi(z[None], 1, 1, length)
        assert np.array_equal(db[0], rb) and np.array_equal(ds[0], rs), length


def t_noise():
    B, T = 4, 20000
    out = np.zeros((B, T, 2), np.float32)
    sig = np.array([1.0, 2.0, 0.5, 1.0], np.float32)
    ok(L.wf_awgn_fill(ptr(out), B, T, ptr(sig), None, 1234, 0, None, None))
    for b in range(B):
        x = out[b].ravel() / sig[b]
        print(f"  noise frame {b}: mean {x.mean():+.4f} var {x.var():.4f} kurt {np.mean(x**4):.3f}")
        assert abs(x.mean()) < 0.03 and abs(x.var() - 1) < 0.03 and abs(np.mean(x ** 4) - 3) < 0.15
    # sharding independence: frames 2..3 generated alone with first_frame=2
    out2 = np.zeros((2, T, 2), np.float32)
    ok(L.wf_awgn_fill(ptr(out2), 2, T, ptr(sig[2:].copy()), None, 1234, 2, None, None))
    assert np.array_equal(out2, out[2:])
    # fused modulate+noise == modulate then add
    sy = o.fsm_encode(o.TRELLISES["SOQPSKTrellis4x2DiffEncoded"], G("g4_modulated.npz")["bits"])[None]
    s1 = modulate(sy, [0.25], o.pulse_soqpsk_tg(8), 8, 0, derot=-np.pi / 4)
    T1 = s1.shape[1]
    sg = np.array([0.7], np.float32)
    buf = np.ascontiguousarray(s1.copy())
    ok(L.wf_awgn_add(ptr(buf), 0, 1, T1, T1, ptr(sg), None, 99, 5, None, None))
    s2 = modulate(sy, [0.25], o.pulse_soqpsk_tg(8), 8, 0, derot=-np.pi / 4, sigma=sg, seed=99, first=5)
    assert np.max(np.abs(buf - s2)) < 1e-6, np.max(np.abs(buf - s2))
    # same identity on every fused kernel shape: the sps-8 fast path with aligned (TG, off 32) and
    # late-starting lanes (multi-h off 12, MIL off 4), a multi-segment frame, the generic kernel
    # (sps 10, float64), with enough pairs that the 2^-15 tail-refinement path is exercised
    rng = np.random.default_rng(11)
    cases = [("tg", 4, 4096, [0.25], o.pulse_soqpsk_tg(8), 8, 0, (-2, 0, 2)),
             ("multih", 4, 2048, [4 / 16, 5 / 16], o.pulse_multih(8), 8, 0, (-3, -1, 1, 3)),
             ("mil", 4, 3000, [0.25], o.pulse_soqpsk_mil(8), 8, 0, (-2, 0, 2)),
             ("pcmfm", 4, 4096, [0.7], o.pulse_pcmfm(8, 4), 8, 0, (-1, 1)),
             ("tg long", 2, 11001, [0.25], o.pulse_soqpsk_tg(8), 8, 0, (-2, 0, 2)),
             ("three-h", 2, 2051, [4 / 16, 5 / 16, 7 / 16], o.pulse_multih(8), 8, 0, (-3, -1, 1, 3)),
             ("tg sps10", 2, 3001, [0.25], o.pulse_soqpsk_tg(10), 10, 0, (-2, 0, 2)),
             ("tg f64", 2, 3001, [0.25], o.pulse_soqpsk_tg(8), 8, 1, (-2, 0, 2))]
    for lab, B, S, h, pulse, sps, dtype, alpha in cases:
        sy = rng.choice(np.array(alpha, np.int8), size=(B, S)).astype(np.int8)
        sg = np.linspace(0.5, 1.5, B).astype(np.float32)
        clean = modulate(sy, h, pulse, sps, dtype)
        Tn = clean.shape[1]
        buf = np.ascontiguousarray(clean.copy())
        ok(L.wf_awgn_add(ptr(buf), dtype, B, Tn, Tn, ptr(sg), None, 77, 3, None, None))
        fused = modulate(sy, h, pulse, sps, dtype, sigma=sg, seed=77, first=3)
        err = np.max(np.abs(buf - fused))
        nz = (buf - clean) / sg[:, None]
        print(f"  fused vs separate {lab}: max diff {err:.2e}; noise var {nz.real.var():.4f} max |n| {np.abs(nz).max():.2f}")
        assert err < 2e-6, (lab, err)
        if dtype == 0 and sps == 8:
            ref = np.stack([o.cpm_modulate_closed_form(sy[b], h if len(h) > 1 else h[0], pulse, sps) for b in range(B)])
            assert np.max(np.abs(clean - ref)) < 1e-5, (lab, np.max(np.abs(clean - ref)))


def t_stream():
    g = G("g5_config1.npz")
    z = np.ascontiguousarray(g["PAM_z"][:301].astype(np.complex128))
    for length in (2, 3, 4):
        rb, rs = o.detect_soqpsk(z, length, True)
        st = np.zeros((1, 5 + 8 * length), np.float64)
        got = []
        for a, b in ((0, 101), (101, 102), (102, 301)):
            zz = np.ascontiguousarray(z[a:b][None])
            db = np.zeros((1, b - a), np.uint8)
            ok(L.wf_viterbi_soqpsk(ptr(zz), 1, 1, b - a, b - a, 1, length, ptr(db), None, None, 0, None, 0, None, None,
                                   ptr(st), None))
            got.append(db[0])
        assert np.array_equal(np.concatenate(got), rb), length
    # frequency_modulate / phase_modulate against the oracle's sequential accumulator
    rng = np.random.default_rng(3)
    f = rng.normal(size=(2, 3000)) * 0.3
    out = np.zeros((2, 3000), np.complex128)
    ok(L.wf_frequency_modulate(ptr(f), 2, 3000, 3000, 8, 0.4, 1, ptr(out), 3000, None))
    for b in range(2):
        revs, ph = 0.0, np.zeros(3000)
        for i in range(3000):
            revs = (revs + f[b, i]) % 8
            ph[i] = revs * 2 * np.pi / 8 + 0.4
        assert np.max(np.abs(out[b] - np.exp(1j * ph))) < 1e-11
    out32 = np.zeros(3000, np.complex64)
    ok(L.wf_phase_modulate(ptr(f[0].copy()), 3000, 1.7, 0, ptr(out32), None))
    assert np.max(np.abs(out32 - np.exp(1j * 1.7 * f[0]))) < 1e-6


def t_lean():
    """counter-only path (no decisions requested): many frames, ragged sizes, per-frame points"""
    rng = np.random.default_rng(5)
    for B, n_calls, S, dtype in ((70, 101, 99, 0), (33, 37, 40, 1), (5, 64, 62, 0), (64, 8, 6, 0), (40, 3, 1, 1)):
        z = (rng.normal(size=(B, n_calls, 3)) + 1j * rng.normal(size=(B, n_calls, 3)))
        zz = np.ascontiguousarray(z.astype(np.complex64 if dtype == 0 else np.complex128))
        db, _ = o.detect_soqpsk_fast(zz.astype(np.complex128), True)
        tx = rng.integers(0, 2, size=(B, S)).astype(np.uint8)
        n_cmp = max(0, min(S, n_calls - 2))
        pts = (np.arange(B) % 3).astype(np.int32)
        want = np.zeros((3, 3), np.int64)
        for b in range(B):
            want[pts[b], 0] += int(np.sum(db[b, 2:2 + n_cmp] != tx[b, :n_cmp]))
            want[pts[b], 2] += n_cmp
        pk = np.packbits(tx, axis=1, bitorder="little")
        pk = np.ascontiguousarray(np.concatenate((pk, np.zeros((B, (-pk.shape[1]) % 4), np.uint8)), axis=1)).view(np.uint32)
        cnt = np.zeros((3, 3), np.int64)
        ok(L.wf_viterbi_soqpsk(ptr(zz), dtype, B, n_calls, n_calls, 1, 2, None, None, ptr(pk), pk.shape[1], None, S,
                               ptr(pts), ptr(cnt), None, None))
        assert np.array_equal(cnt, want), (B, n_calls, S, dtype, cnt, want)


def t_rawbank():
    """bank without combine + wf_viterbi_soqpsk_combined == bank with combine + wf_viterbi_soqpsk"""
    from waveforms_b200 import mfbank as _mf
    rng = np.random.default_rng(9)
    sps = 8
    taps, comb = _mf.pam_taps(o.pulse_soqpsk_tg(sps), 0.25, sps)
    taps, comb = np.asarray(taps), np.asarray(comb)
    for B, S, dtype in ((37, 300, 0), (3, 77, 0), (34, 129, 1)):
        T = (S + 1) * sps
        first, n_calls = _mf.symbol_instants(T, sps, 0, 2)
        rx = (rng.normal(size=(B, T)) + 1j * rng.normal(size=(B, T)))
        ld = (n_calls + 3) & ~3
        cdt = np.complex64 if dtype == 0 else np.complex128
        rx = np.ascontiguousarray(rx.astype(cdt))
        tr, trp = darr(taps.real.ravel())
        cr, crp = darr(comb.real.ravel())
        ci, cip = darr(comb.imag.ravel())
        z = np.zeros((B, ld, 3), cdt)
        y = np.zeros((B, ld, 2), cdt)
        c = (taps.shape[1] - 1) // 2
        ok(L.wf_mf_bank(ptr(rx), dtype, B, T, T, trp, None, 2, taps.shape[1], c, crp, cip, 3, sps, first, n_calls, ptr(z), ld, None))
        ok(L.wf_mf_bank(ptr(rx), dtype, B, T, T, trp, None, 2, taps.shape[1], c, None, None, 2, sps, first, n_calls, ptr(y), ld, None))
        tx = rng.integers(0, 2, size=(B, S)).astype(np.uint8)
        pk = np.packbits(tx, axis=1, bitorder="little")
        pk = np.ascontiguousarray(np.concatenate((pk, np.zeros((B, (-pk.shape[1]) % 4), np.uint8)), axis=1)).view(np.uint32)
        pts = (np.arange(B) % 2).astype(np.int32)
        ca, cb = np.zeros((2, 3), np.int64), np.zeros((2, 3), np.int64)
        ok(L.wf_viterbi_soqpsk(ptr(z), dtype, B, n_calls, ld, 1, 2, None, None, ptr(pk), pk.shape[1], None, S, ptr(pts), ptr(ca),
                               None, None))
        ok(L.wf_viterbi_soqpsk_combined(ptr(y), dtype, B, n_calls, ld, 2, crp, cip, 1, ptr(pk), pk.shape[1], S, ptr(pts),
                                        ptr(cb), None))
        print(f"  raw bank B={B} S={S} dtype={dtype}: counters {ca.tolist()} vs {cb.tolist()}")
        assert np.array_equal(ca, cb) and ca[:, 2].sum() == B * max(0, min(S, n_calls - 2))


def t_modbank():
    """wf_cpm_modulate_bank (block kernel and the two warp-autonomous kernels) == wf_cpm_modulate + wf_mf_bank, bit for bit"""
    from waveforms_b200 import mfbank as _mf
    rng = np.random.default_rng(4)
    sps = 8
    pulse = o.pulse_soqpsk_tg(sps)
    taps, _comb = _mf.pam_taps(pulse, 0.25, sps)
    taps = np.asarray(taps)
    tr, trp = darr(taps.real.ravel())
    hv, hp = darr([0.25])
    pv, pp = darr(pulse)
    c = (taps.shape[1] - 1) // 2
    for B, S in ((3, 40), (5, 200), (2, 520), (3, 1031), (1, 2100)):
        T = (S + 1) * sps
        first, n_calls = _mf.symbol_instants(T, sps, 0, 2)
        ld = (n_calls + 3) & ~3
        syms = rng.choice(np.array([-2, 0, 2], np.int8), size=(B, S)).astype(np.int8)
        sg = np.linspace(0.0, 1.5, B).astype(np.float32)
        pts = (np.arange(B) % 2).astype(np.int32)
        fid = (1000 + 7 * np.arange(B)).astype(np.uint64)
        ldrx = T + (T & 1)
        rx = np.zeros((B, ldrx), np.complex64)
        ok(L.wf_cpm_modulate(ptr(syms), B, S, hp, 1, pp, pv.size, sps, np.pi / 4, -np.pi / 4, 0, ptr(rx), ldrx, ptr(sg), ptr(pts), 77, 0,
                             ptr(fid), None))
        y = np.full((B, ld, 2), np.nan, np.complex64)
        ok(L.wf_mf_bank(ptr(rx), 0, B, T, ldrx, trp, None, 2, taps.shape[1], c, None, None, 2, sps, first, n_calls, ptr(y), ld, None))
        assert L.wf_cpm_modulate_bank_supported(pv.size, sps, 2, taps.shape[1], c, first, S, n_calls) == 1
        for tune in (0, 1):
            z = np.full((B, ld, 2), np.nan, np.complex64)
            ok(L.wf_cpm_modulate_bank(ptr(syms), B, S, hp, 1, pp, pv.size, sps, np.pi / 4, -np.pi / 4, ptr(sg), ptr(pts), 77, 0, ptr(fid),
                                      trp, 2, taps.shape[1], c, first, n_calls, ptr(z), ld, tune, None))
            same = np.array_equal(z[:, :n_calls].view(np.uint32), y[:, :n_calls].view(np.uint32))
            bad = int(np.sum(z[:, :n_calls].view(np.uint32) != y[:, :n_calls].view(np.uint32)))
            print(f"  B={B} S={S} tune={tune}: fused bank outputs identical to the two-kernel path: {same} ({bad} words differ)")
            assert same, (B, S, tune)


def t_linkcall():
    """wf_link_sim_fused == wf_prbs_encode + wf_cpm_modulate_bank + wf_viterbi_soqpsk_combined (counters), and
    wf_soqpsk_detect_static == wf_mf_bank + wf_viterbi_soqpsk (decisions + counters), twice in a row"""
    from waveforms_b200 import mfbank as _mf
    sps, deg = 8, 15
    pulse = o.pulse_soqpsk_tg(sps)
    taps, comb = _mf.pam_taps(pulse, 0.25, sps)
    taps, comb = np.asarray(taps), np.asarray(comb)
    tr, trp = darr(taps.real.ravel())
    cr, crp = darr(comb.real.ravel())
    ci, cip = darr(comb.imag.ravel())
    hv, hp = darr([0.25])
    pv, pp = darr(pulse)
    c = (taps.shape[1] - 1) // 2
    mask, per = o.pn_mask(deg), (1 << deg) - 1
    for B, N in ((5, 200), (3, 520), (4, 96)):
        T = (N + 1) * sps
        first, n_calls = _mf.symbol_instants(T, sps, 0, 2)
        ld, ldw = (n_calls + 3) & ~3, (N + 31) // 32
        sg = np.linspace(0.3, 1.2, B).astype(np.float32)
        pts = (np.arange(B) % 2).astype(np.int32)
        fid = (50 + 3 * np.arange(B)).astype(np.uint64)
        boff = ((fid * N) % per).astype(np.uint64)
        packed = np.zeros((B, ldw), np.uint32)
        syms = np.zeros((B, N), np.int8)
        z = np.zeros((B, ld, 2), np.complex64)
        want = np.zeros((2, 3), np.int64)
        ok(L.wf_prbs_encode(mask, deg, per, per, ptr(boff), 0, B, N, 0, None, None, ptr(packed), ldw, ptr(syms), None))
        ok(L.wf_cpm_modulate_bank(ptr(syms), B, N, hp, 1, pp, pv.size, sps, np.pi / 4, -np.pi / 4, ptr(sg), ptr(pts), 9, 0, ptr(fid),
                                  trp, 2, taps.shape[1], c, first, n_calls, ptr(z), ld, 1, None))
        ok(L.wf_viterbi_soqpsk_combined(ptr(z), 0, B, n_calls, ld, 2, crp, cip, 1, ptr(packed), ldw, N, ptr(pts), ptr(want), None))
        cfg = _ffi.LinkCfg(mask=mask, init_state=per, period=per, degree=deg, encoder=0, n_bits=N, mod_index=hp, pulse=pp, n_h=1,
                           K=pv.size, sps=sps, tune=1, initial_phase=np.pi / 4, derotate=-np.pi / 4, seed=9, taps_re=trp, n_filt=2,
                           n_taps=taps.shape[1], center=c, diff=1, first_instant=first, n_calls=n_calls, combine_re=crp,
                           combine_im=cip)
        need = L.wf_link_sim_workspace_bytes(C.byref(cfg), B)
        ws = np.zeros(need + 256, np.uint8)
        a = (-ws.ctypes.data) % 256
        got = np.zeros((2, 3), np.int64)
        ok(L.wf_link_sim_fused(C.byref(cfg), B, 0, ptr(fid), ptr(boff), ptr(sg), ptr(pts), C.c_void_p(ws.ctypes.data + a), need,
                               ptr(got), None))
        assert np.array_equal(got, want) and want[:, 2].sum() > 0, (B, N, got, want)
        assert L.wf_link_sim_fused(C.byref(cfg), B, 0, ptr(fid), ptr(boff), ptr(sg), ptr(pts), C.c_void_p(ws.ctypes.data + a),
                                   need - 1, ptr(got), None) == -1
        # static-scratch detector on the noisy frames of the un-fused chain
        ldrx = T + (T & 1)
        rx = np.zeros((B, ldrx), np.complex64)
        ok(L.wf_cpm_modulate(ptr(syms), B, N, hp, 1, pp, pv.size, sps, np.pi / 4, -np.pi / 4, 0, ptr(rx), ldrx, ptr(sg), ptr(pts), 9, 0,
                             ptr(fid), None))
        z3 = np.zeros((B, n_calls, 3), np.complex64)
        wc, wb, wsy = np.zeros((2, 3), np.int64), np.zeros((B, n_calls), np.uint8), np.zeros((B, n_calls), np.int8)
        ok(L.wf_mf_bank(ptr(rx), 0, B, T, ldrx, trp, None, 2, taps.shape[1], c, crp, cip, 3, sps, first, n_calls, ptr(z3), n_calls, None))
        ok(L.wf_viterbi_soqpsk(ptr(z3), 0, B, n_calls, n_calls, 1, 2, ptr(wb), ptr(wsy), ptr(packed), ldw, ptr(syms), N, ptr(pts),
                               ptr(wc), None, None))
        hc, hb, hsy = np.zeros_like(wc), np.zeros_like(wb), np.zeros_like(wsy)
        ok(L.wf_soqpsk_detect_static(ptr(rx), B, T, ldrx, trp, None, 2, taps.shape[1], c, crp, cip, sps, first, n_calls, 1, 2, ptr(hb),
                                     ptr(hsy), ptr(packed), ldw, ptr(syms), N, ptr(pts), ptr(hc), None))
        assert np.array_equal(hc, wc) and np.array_equal(hb, wb) and np.array_equal(hsy, wsy), (B, N, hc, wc)
        assert np.array_equal(wc[:, [0, 2]], want[:, [0, 2]])  # and both chains count the same frames identically
        print(f"  B={B} N={N}: one-call link counters {got.tolist()} == three calls; static-scratch detector == two calls")
    lim = L.wf_static_scratch_max_calls()
    assert L.wf_soqpsk_detect_static(None, lim // 100 + 1, 808, 808, trp, None, 2, taps.shape[1], c, crp, cip, sps, 0, 100, 1, 2, None,
                                     None, None, 0, None, 0, None, None, None) == -1
    assert b"static scratch" in L.wf_last_error_string()


def t_diag():
    """csrc/diag.cu (Welch PSD, eye traces, phase tree, timing metric) vs oracle/viz_ref.py and golden G12"""
    from oracle import viz_ref as vr
    rng = np.random.default_rng(8)
    g = G("g12_phase_tree.npz")
    pulse = o.pulse_soqpsk_tg(8)
    sy = o.fsm_encode(o.TRELLISES["SOQPSKTrellis4x2DiffEncoded"], rng.integers(0, 2, 700).astype(np.uint8))
    s = o.cpm_modulate_closed_form(sy, 0.25, pulse, 8)
    sig = np.ascontiguousarray(np.stack([s, s * np.exp(0.3j) + 0.05 * (rng.normal(size=s.size) + 1j * rng.normal(size=s.size))]).astype(np.complex64))
    B, T = sig.shape
    for nfft in (64, 1024):
        out = np.zeros((B, nfft), np.float32)
        ok(L.wf_welch_psd(ptr(sig), B, T, T, nfft, float(np.sqrt(2)), ptr(out), None))
        for b in range(B):
            _f, ref = vr.psd(sig[b].astype(np.complex128), 8, 2, nfft)
            rel = np.max(np.abs(out[b] - ref) / (ref + 1e-9 * ref.max()))
            assert rel < 2e-3, (nfft, b, rel)
    span = 32
    n_tr = (T - 1) // span
    eye = np.zeros((B, n_tr, span + 1), np.complex64)
    ok(L.wf_eye_traces(ptr(sig), B, n_tr * span + 1, T, span, ptr(eye), None))
    for b in range(B):
        _t, ref = vr.eye_traces(np.arange(T) / 8.0, sig[b], 8, 4)
        assert np.array_equal(eye[b], ref)
    for span, off in ((32, None), (24, 0.25), (100, None)):
        nc = T // span
        ph = np.zeros((B, nc, span), np.float32)
        ok(L.wf_phase_tree(ptr(sig), B, T, T, span, int(off is not None), float(off or 0.0), ptr(ph), None))
        for b in range(B):
            phase = np.angle(sig[b].astype(np.complex128))
            for c in range(nc):
                sl = np.unwrap(phase[c * span:(c + 1) * span])
                ref = sl - (off if off is not None else sl[0])
                assert np.max(np.abs(ph[b, c] - ref)) < 2e-5, (span, b, c)
    for lab, hh, pf in (("pcmfm", [0.7], o.pulse_pcmfm(8, 4)), ("multih", [4 / 16, 5 / 16], o.pulse_multih(8))):
        syms = g[f"{lab}_symbols"]
        length = pf.size // 8
        plotted = np.zeros(syms.shape[0] * length * 8 + 1, np.complex64)
        for i, q in enumerate(syms):
            plotted[i * length * 8:(i + 1) * length * 8] = o.cpm_modulate_closed_form(q, np.array(hh), pf, 8)[length * 8 - 1:2 * length * 8 - 1]
        nc = plotted.size // (length * 8)
        ph = np.zeros((1, nc, length * 8), np.float32)
        ok(L.wf_phase_tree(ptr(plotted), 1, plotted.size, plotted.size, length * 8, 0, 0.0, ptr(ph), None))
        assert np.max(np.abs(ph[0] - g[f"{lab}_phases"])) < 1e-4, lab
    for sps in (8, 10, 3):
        met = np.zeros((B, sps), np.float32)
        ok(L.wf_timing_metric(ptr(sig), B, T, T, sps, ptr(met), None))
        for b in range(B):
            assert np.max(np.abs(met[b] - vr.timing_metric(sig[b].astype(np.complex128), sps))) < 1e-5


def t_cpm():
    """generic CPM detector (PCM/FM, multi-h) vs oracle/cpm_generic.py, bank + trellis"""
    from oracle.cpm_generic import CpmTrellis
    rng = np.random.default_rng(11)
    cases = (("pcmfm", o.pulse_pcmfm(8, 4), [7], 10, 2, 2, 512), ("pcmfm1", o.pulse_pcmfm(8, 4), [7], 10, 2, 1, 256),
             ("multih", o.pulse_multih(8), [4, 5], 16, 4, 2, 600), ("multih1", o.pulse_multih(8), [4, 5], 16, 4, 1, 200))
    for name, pulse, hnum, p, M, Lp, nbits in cases:
        tr = CpmTrellis(pulse, hnum, p, M, 8, Lp)
        B = 3
        bits = rng.integers(0, 2, (B, nbits)).astype(np.uint8)
        syms = np.stack([o.map_pcmfm(b) if M == 2 else o.map_multih(b) for b in bits])
        S = syms.shape[1]
        bps = 1 if M == 2 else 2
        sg = o.sigma_from_ebn0_db(5.0, 8, bps)
        rx = np.stack([o.cpm_modulate_closed_form(sy, np.array(hnum) / p, pulse, 8) * np.exp(-1j * np.pi / 4) for sy in syms])
        rx = rx + sg * (rng.normal(size=rx.shape) + 1j * rng.normal(size=rx.shape))
        taps = tr.bank_taps()
        for dtype in (1, 0):
            z = mf_center(rx, taps, 8, S, dtype, tr.first_instant, 7)
            zref = np.stack([tr.bank(r, S) for r in rx])
            err = np.max(np.abs(z - zref))
            assert err < (1e-9 if dtype else 2e-4), (name, dtype, err)
            zz = np.ascontiguousarray(z)
            dec = np.zeros((B, S), np.int8)
            cnt = np.zeros((1, 3), np.int64)
            hn = np.array(hnum, np.int32)
            ok(L.wf_viterbi_cpm(ptr(zz), dtype, B, S, S, M, Lp, p, ptr(hn), len(hnum), 0, ptr(dec), ptr(np.ascontiguousarray(syms)), S, 4, 4,
                                None, ptr(cnt), None))
            ref = np.stack([2 * tr.detect(zref[b]).astype(np.int64) - (M - 1) for b in range(B)])
            nd = int(np.sum(dec != ref))
            se = int(np.sum(ref[:, 4:S - 4] != syms[:, 4:S - 4]))
            print(f"  {name} dtype={dtype}: bank err {err:.1e}, decisions differing from oracle {nd}, symbol errors {cnt[0,1]} (oracle {se})")
            if dtype == 1:
                assert nd == 0 and cnt[0, 1] == se and cnt[0, 2] == B * (S - 8) * bps
            else:
                assert nd <= 2


def t_cpmfused():
    """wf_cpm_detect (bank fused into the detector, decision bits + block traceback) vs oracle/cpm_generic.py"""
    from oracle.cpm_generic import CpmTrellis
    rng = np.random.default_rng(12)
    cases = (("pcmfm", o.pulse_pcmfm(8, 4), [7], 10, 2), ("multih", o.pulse_multih(8), [4, 5], 16, 4))
    for name, pulse, hnum, p, M in cases:
        tr = CpmTrellis(pulse, hnum, p, M, 8, 2)
        bps = 1 if M == 2 else 2
        taps = tr.bank_taps()
        tre, tim = np.ascontiguousarray(taps.real.ravel()), np.ascontiguousarray(taps.imag.ravel())
        hn = np.array(hnum, np.int32)
        for S, depth, ebn0, tune in ((200, 0, 5.0, 0), (67, 5, 3.0, 0), (33, 16, 4.0, 2), (130, 32, 2.0, 4), (7, 0, 6.0, 0), (1, 0, 6.0, 0),
                                     (96, 8, 1.0, 0)):
            B = 3 if name == "multih" else 35   # PCM/FM: more than one warp, a ragged last warp
            bits = rng.integers(0, 2, (B, S * bps)).astype(np.uint8)
            syms = np.ascontiguousarray(np.stack([o.map_pcmfm(b) if M == 2 else o.map_multih(b) for b in bits]))
            sg = o.sigma_from_ebn0_db(ebn0, 8, bps)
            rx = np.stack([o.cpm_modulate_closed_form(sy, np.array(hnum) / p, pulse, 8) * np.exp(-1j * np.pi / 4) for sy in syms])
            rx = (rx + sg * (rng.normal(size=rx.shape) + 1j * rng.normal(size=rx.shape))).astype(np.complex64)
            T = rx.shape[1]
            assert T == (S + 1) * 8
            sup = L.wf_cpm_detect_supported(0, T, T, ptr(tre), ptr(tim), 8, tr.first_instant, S, M, 2, p, ptr(hn), len(hnum), depth)
            assert sup == 1, (name, S, depth)
            dec = np.full((B, S), 99, np.int8)
            cnt = np.zeros((1, 3), np.int64)
            skip = 4 if S > 8 else 0
            ok(L.wf_cpm_detect(ptr(rx), 0, B, T, T, ptr(tre), ptr(tim), 8, tr.first_instant, S, M, 2, p, ptr(hn), len(hnum), depth,
                               ptr(dec), ptr(syms), S, skip, skip, None, ptr(cnt), tune, None))
            D = depth or 16
            nd = 0
            for b in range(B):
                ref, mg = tr.detect(tr.bank(rx[b].astype(np.complex128), S), depth=D, return_margin=True)
                ref = 2 * ref.astype(np.int64) - (M - 1)
                rel = mg[:, 0] / np.maximum(mg[:, 1], 1.0)
                for k in np.flatnonzero(dec[b] != ref):
                    nd += 1
                    m = rel[k:k + 2 * D].min()
                    assert m < 1e-5, (name, S, depth, b, k, m)
            se = int(np.sum(dec[:, skip:S - skip] != syms[:, skip:S - skip]))
            be = int(sum(bin(int(x)).count("1") for x in (((dec[:, skip:S - skip].astype(int) + M - 1) >> 1) ^
                                                         ((syms[:, skip:S - skip].astype(int) + M - 1) >> 1)).ravel()))
            assert cnt[0].tolist() == [be, se, B * max(0, S - 2 * skip) * bps], (name, S, depth, cnt, be, se)
            print(f"  {name} S={S} depth={depth} tune={tune}: {nd} decisions differ from the float64 oracle (all near-ties), "
                  f"symbol errors {se}")
        # unsupported shapes are refused, not run
        assert L.wf_cpm_detect_supported(1, 808, 808, ptr(tre), ptr(tim), 8, tr.first_instant, 100, M, 2, p, ptr(hn), len(hnum), 0) == 0
        assert L.wf_cpm_detect_supported(0, 808, 808, ptr(tre), ptr(tim), 8, 0, 100, M, 2, p, ptr(hn), len(hnum), 0) == 0
        bad = tre.copy(); bad[3] += 0.25
        assert L.wf_cpm_detect_supported(0, 808, 808, ptr(bad), ptr(tim), 8, tr.first_instant, 100, M, 2, p, ptr(hn), len(hnum), 0) == 0


TESTS = {"lfsr": t_lfsr, "prbs": t_prbs, "mod": t_mod, "modedge": t_modedge, "mf": t_mf, "vit": t_vit, "noise": t_noise, "stream": t_stream, "lean": t_lean, "rawbank": t_rawbank, "cpm": t_cpm, "cpmfused": t_cpmfused, "modbank": t_modbank, "diag": t_diag, "linkcall": t_linkcall}

if __name__ == "__main__":
    names = sys.argv[1:] or list(TESTS)
    for n in names:
        t0 = time.time()
        print(f"[{n}]")
        TESTS[n]()
        print(f"[{n}] ok in {time.time() - t0:.1f}s")
