"""GPU parity of the multibeam coincidencer through the C ABI: the matching
kernel against the unmodified reference kernel run on the same device buffers (oracle/_ref) and
against oracle/coincidence.py (bit-exact), and the whole program on synthetic beams with
interference common to all beams and a pulsar in one."""
import ctypes as C
import os

import numpy as np
import pytest

import peasoup_b200 as pb
from synth import write_sigproc

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


@pytest.mark.parametrize("narrays,size,beam_thresh", [(1, 1000, 1), (4, 100003, 2), (13, 1 << 20, 4), (7, 5, 7)])
def test_coincidencer_kernel_bit_exact(lib, narrays, size, beam_thresh):
    from oracle import coincidence as K
    rng = np.random.default_rng(size)
    arrs = [rng.normal(0, 2, size).astype(np.float32) for _ in range(narrays)]
    d = [torch.from_numpy(a).cuda() for a in arrs]
    table = torch.tensor([t.data_ptr() for t in d], dtype=torch.int64, device="cuda")
    out = torch.full((size,), -1.0, dtype=torch.float32, device="cuda")
    pb.check(lib.ps_coincidencer(table.data_ptr(), out.data_ptr(), narrays, size, 1.5, beam_thresh, None))
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), K.match(arrs, 1.5, beam_thresh))


@pytest.mark.parametrize("narrays,size,beam_thresh", [(1, 1000, 1), (4, 100003, 2), (13, 1 << 20, 4), (7, 5, 7), (6, 40000, 4)])
def test_coincidencer_kernel_equals_reference_kernel(lib, ref, narrays, size, beam_thresh):
    """Pin against the UNMODIFIED reference kernel run here: coincidence_kernel / device_coincidencer
    (kernels.cu:1073-1100, forwarded by oracle/ref_harness.cu).  The reference kernel has no bounds
    check (whole blocks of 512 threads run past `size`), so its buffers are padded to a multiple of
    512; the first `size` mask values must be identical to ours, including values equal to the
    threshold (strict >) and NaNs."""
    rng = np.random.default_rng(size + narrays)
    padded = (size + 511) // 512 * 512
    arrs = []
    for i in range(narrays):
        a = rng.normal(0, 2, padded).astype(np.float32)
        a[rng.integers(0, size, max(1, size // 50))] = np.float32(1.5)   # exactly at the threshold
        if size > 100:
            a[rng.integers(0, size, 3)] = np.nan
        arrs.append(a)
    d = [torch.from_numpy(a).cuda() for a in arrs]
    table = torch.tensor([t.data_ptr() for t in d], dtype=torch.int64, device="cuda")
    ours = torch.full((padded,), -1.0, dtype=torch.float32, device="cuda")
    theirs = torch.full((padded,), -1.0, dtype=torch.float32, device="cuda")
    pb.check(lib.ps_coincidencer(table.data_ptr(), ours.data_ptr(), narrays, size, 1.5, beam_thresh, None))
    ref.ref_coincidencer.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_size_t, C.c_float, C.c_int]
    assert ref.ref_coincidencer(table.data_ptr(), theirs.data_ptr(), narrays, size, 1.5, beam_thresh) == 0, ref.ref_last_error()
    torch.cuda.synchronize()
    assert torch.equal(ours[:size], theirs[:size])
    from oracle import coincidence as K                                   # ... and so must the oracle's restatement
    assert np.array_equal(theirs[:size].cpu().numpy(), K.match([a[:size] for a in arrs], 1.5, beam_thresh))
    assert torch.all(ours[size:] == -1.0)                                # ours never writes past `size`
    assert set(np.unique(ours[:size].cpu().numpy())) <= {0.0, 1.0}


def test_coincidencer_rejects_empty(lib):
    assert lib.ps_coincidencer(None, None, 0, 10, 1.0, 1, None) == -2


def make_beams(tmp_path, nbeams=6, nsamps=40000, nchans=16, tsamp=0.00032):
    rng = np.random.default_rng(7)
    t = np.arange(nsamps) * tsamp
    rfi_tone = 6.0 * np.sin(2 * np.pi * 50.0 * t)                       # mains birdie in every beam
    spikes = np.zeros(nsamps)
    spikes[[5000, 5001, 17000, 31234]] = 90.0                           # broadband impulses in every beam
    pulsar = 14.0 * ((t / 0.0123) % 1.0 < 0.05)                          # celestial: beam 2 only
    files = []
    for b in range(nbeams):
        x = 100 + rng.normal(0, 6, (nsamps, nchans)) + (rfi_tone + spikes)[:, None]
        if b == 2:
            x = x + pulsar[:, None]
        path = os.path.join(tmp_path, "beam%02d.fil" % b)
        write_sigproc(path, np.clip(np.rint(x), 0, 255).astype(np.uint8), tsamp, 1510.0, -1.0)
        files.append(path)
    return files, tsamp


def test_coincidence_program_matches_oracle(lib, oracle, tmp_path):
    from oracle import coincidence as K
    files, tsamp = make_beams(str(tmp_path))
    n = 40000
    samp_file, spec_file = os.path.join(tmp_path, "rfi.eb_mask"), os.path.join(tmp_path, "birdies.txt")
    prm = pb.CoincidenceParams(4.0, 4, samp_file.encode(), spec_file.encode())
    arr = (C.c_char_p * len(files))(*[f.encode() for f in files])
    samp = np.full(n, -1, np.float32)
    spec = np.full(n // 2 + 1, -1, np.float32)
    got = lib.ps_coincidence_files(arr, len(files), C.byref(prm), samp.ctypes.data, n, spec.ctypes.data, n // 2 + 1)
    assert got == n, lib.ps_last_error()
    # oracle on the same files
    tims, specs = [], []
    for f in files:
        hdr, data = oracle.read_sigproc(f)
        trial, _ = oracle.dedisperse(data, hdr, np.zeros(1, np.float32))
        assert trial.shape == (1, n)
        t_, p_, bw = K.whiten_beam(trial[0], np.float32(hdr["tsamp"]))
        tims.append(t_)
        specs.append(p_)
    for mask, arrays in ((samp, tims), (spec, specs)):
        want = K.match(arrays, 4.0, 4)
        # a value within 2e-3 of the threshold may fall on either side (FFT rounding)
        near = np.min([np.abs(a - np.float32(4.0)) for a in arrays], axis=0) < 2e-3
        bad = (mask != want) & ~near
        assert not bad.any(), (int(bad.sum()), np.flatnonzero(bad)[:5])
        assert set(np.unique(mask)) <= {0.0, 1.0}
    # the interference is flagged, the single-beam pulsar is not
    assert all(samp[i] == 0 for i in (5000, 5001, 17000, 31234)) and samp.mean() > 0.999
    k50 = int(round(50.0 * n * tsamp))
    assert spec[k50] == 0
    k_psr = int(round(n * tsamp / 0.0123))
    assert spec[k_psr - 1:k_psr + 2].min() == 1
    # files as the reference writes them
    assert open(samp_file).read() == K.samp_mask_text(samp)
    assert open(spec_file).read() == K.birdie_text(K.birdies_from_mask(spec, bw))
    assert len(open(spec_file).read().splitlines()) >= 1


def test_coincidence_program_errors(lib, tmp_path):
    prm = pb.CoincidenceParams(4.0, 4, None, None)
    arr = (C.c_char_p * 1)(b"/nonexistent.fil")
    assert lib.ps_coincidence_files(arr, 1, C.byref(prm), None, 0, None, 0) < 0
    assert b"could not be opened" in lib.ps_last_error()
