"""Developer diagnostic (GPU box): where in the frame do the high-SNR bit errors of the SOQPSK-TG
PAM link sit, and do they depend on the noise generator?  Runs the chain stage by stage through
the public API with (a) the library's Philox noise and (b) torch.randn noise of the same sigma."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from waveforms_b200 import mfbank as mf  # noqa: E402
from waveforms_b200.cpm import modulate as wmod  # noqa: E402
from waveforms_b200.cpm.soqpsk import freq_pulse_soqpsk_tg  # noqa: E402
from waveforms_b200.linksim import LinkConfig, LinkSimulator  # noqa: E402
from waveforms_b200.noise import PhiloxRNG, add_awgn, sigma_from_ebn0_db  # noqa: E402
from waveforms_b200.viterbi.algorithm import SOQPSKTrellisDetector  # noqa: E402

dev = torch.device("cuda", 0)
EB = float(sys.argv[1]) if len(sys.argv) > 1 else 14.0
B, N, sps = 16384, 4096, 8
rounds = int(sys.argv[2]) if len(sys.argv) > 2 else 8
sim = LinkSimulator(LinkConfig(n_bits=N, detector="PAM", extra={"combine_in_bank": True}), device=dev, max_batch=B)
lib_ffi = __import__("waveforms_b200._ffi", fromlist=["x"])
sigma = float(sigma_from_ebn0_db(EB, sps))
pulse = freq_pulse_soqpsk_tg(sps)
det = SOQPSKTrellisDetector(2, differantial_encoding=True)
for label in ("philox", "torch.randn"):
    hist = np.zeros(N, np.int64)
    tot = 0
    for r in range(rounds):
        first = 1000000 + r * B
        sg = torch.full((B,), sigma, dtype=torch.float32, device=dev)
        cnt = torch.zeros((1, 3), dtype=torch.int64, device=dev)
        sim.run_batch(first, torch.zeros_like(sg), None, cnt)  # fills sim.syms / sim.packed (sigma 0: result unused)
        clean = wmod.modulate_frames(sim.syms, [0.25], pulse, sps, derotate=-np.pi / 4).contiguous()
        if label == "philox":
            rx = add_awgn(clean, sg, PhiloxRNG(1), first_frame=first)
        else:
            g = torch.Generator(device=dev); g.manual_seed(1234 + r)
            nz = torch.randn((B, clean.shape[1], 2), generator=g, device=dev, dtype=torch.float32) * sigma
            rx = clean + torch.view_as_complex(nz)
        z = mf.pam_bank(rx, pulse, 0.25, sps)
        bits, _ = det.detect(z.contiguous(), return_decisions=True, n_calls=sim.n_calls)
        words = sim.packed.view(torch.int32).cpu().numpy().view(np.uint32)
        tx = np.unpackbits(words.view(np.uint8), axis=1, bitorder="little")[:, :N]
        dec = bits.cpu().numpy()[:, 2:2 + sim.n_compared]
        err = dec != tx[:, :sim.n_compared]
        hist[:sim.n_compared] += err.sum(0)
        tot += err.size
    e = int(hist.sum())
    edge = int(hist[:16].sum() + hist[sim.n_compared - 16:sim.n_compared].sum())
    print(f"{label:12s} Eb/N0 {EB} dB: {e} errors in {tot:.3e} bits (BER {e / tot:.3e}); first/last 16 positions hold {edge}; "
          f"top positions {np.argsort(hist)[-8:][::-1].tolist()} counts {np.sort(hist)[-8:][::-1].tolist()}")
