"""Host-side pieces of the pre-stage (row f1) that need no GPU: DM list generation and max delay
against the reference's golden overview.xml and the oracle."""
import ctypes as C
import os

import numpy as np

from conftest import GOLDEN


def test_dm_list_matches_golden_overview(lib, oracle):
    hdr, _ = oracle.read_sigproc(os.path.join(GOLDEN, "tutorial.fil"))
    ov = oracle.parse_overview_xml(os.path.join(GOLDEN, "overview.xml"))
    out = np.zeros(4096, np.float32)
    n = lib.ps_dedisp_generate_dm_list(0.0, 250.0, hdr["tsamp"], 64.0, hdr["fch1"], hdr["foff"], hdr["nchans"], 1.1,
                                       out.ctypes.data, len(out))
    assert n == 59 == lib.ps_dedisp_generate_dm_list(0.0, 250.0, hdr["tsamp"], 64.0, hdr["fch1"], hdr["foff"], hdr["nchans"], 1.1,
                                                     None, 0)
    assert np.allclose(out[:n], ov["dm_list"], rtol=2e-6, atol=1e-4)
    assert np.allclose(out[:n], oracle.dedisp_dm_list(0, 250, hdr["tsamp"], 64.0, hdr["fch1"], hdr["foff"], hdr["nchans"], 1.1),
                       rtol=1e-6, atol=1e-6)
    assert out[0] == 0.0 and np.all(np.diff(out[:n]) > 0) and out[n - 2] < 250.0 <= out[n - 1]


def test_max_delay_matches_oracle(lib, oracle, tutorial_trials):
    trials, ov = tutorial_trials
    hdr, _ = oracle.read_sigproc(os.path.join(GOLDEN, "tutorial.fil"))
    dms = np.ascontiguousarray(ov["dm_list"], np.float32)
    d = lib.ps_dedisp_max_delay(dms.ctypes.data, len(dms), hdr["nchans"], hdr["tsamp"], hdr["fch1"], hdr["foff"])
    assert d == hdr["nsamples"] - trials.shape[1] == 140


def test_degenerate_dm_range(lib):
    out = np.zeros(8, np.float32)
    assert lib.ps_dedisp_generate_dm_list(5.0, 5.0, 64e-6, 64.0, 1510.0, -1.09, 64, 1.1, out.ctypes.data, 8) == 1 and out[0] == 5.0
