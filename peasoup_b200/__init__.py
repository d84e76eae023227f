"""peasoup_b200 -- ctypes binding of libpeasoup_b200.so (the C ABI in include/peasoup_b200.h).

Python is plumbing only (device memory through torch tensors, process launch, tests and
bench); every computation is a hand-written sm_100a kernel inside the shared library.  There
is no CPU fallback: if the library is missing, ``load()`` raises.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_ROOT = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(_ROOT)
LIB_PATH = os.path.join(_ROOT, "lib", "libpeasoup_b200.so")
_lib = None


class PeasoupError(RuntimeError):
    pass


class SearchParams(C.Structure):
    _fields_ = [("size", C.c_uint), ("nsamps", C.c_uint), ("tsamp", C.c_float), ("nharmonics", C.c_int),
                ("min_snr", C.c_float), ("min_freq", C.c_float), ("max_freq", C.c_float),
                ("boundary_5_freq", C.c_float), ("boundary_25_freq", C.c_float), ("max_peaks", C.c_uint)]


class PipelineParams(C.Structure):
    _fields_ = [("search", SearchParams), ("freq_tol", C.c_float), ("max_harm", C.c_int),
                ("acc_start", C.c_float), ("acc_end", C.c_float), ("acc_tol", C.c_float),
                ("acc_pulse_width", C.c_float), ("cfreq", C.c_float), ("foff", C.c_float)]


class FilterbankParams(C.Structure):
    _fields_ = [("pipe", PipelineParams), ("dm_start", C.c_float), ("dm_end", C.c_float), ("dm_tol", C.c_float),
                ("dm_pulse_width", C.c_float), ("limit", C.c_int), ("nworkers", C.c_int), ("device0", C.c_int),
                ("killfile", C.c_char_p), ("zapfile", C.c_char_p), ("outdir", C.c_char_p), ("npdmp", C.c_int)]


class CoincidenceParams(C.Structure):
    _fields_ = [("threshold", C.c_float), ("beam_threshold", C.c_int), ("samp_outfilename", C.c_char_p),
                ("spec_outfilename", C.c_char_p)]


class FoldResult(C.Structure):
    _fields_ = [("opt_sn", C.c_float), ("sn1", C.c_float), ("sn2", C.c_float), ("opt_period", C.c_double),
                ("opt_width", C.c_int), ("opt_bin", C.c_int), ("opt_shift", C.c_int), ("argmax", C.c_uint),
                ("peak", C.c_float)]


class RawPeak(C.Structure):
    _fields_ = [("acc_idx", C.c_int32), ("nh", C.c_int32), ("bin", C.c_int32), ("snr", C.c_float)]


class Candidate(C.Structure):
    _fields_ = [("dm", C.c_float), ("dm_idx", C.c_int32), ("acc", C.c_float), ("nh", C.c_int32),
                ("snr", C.c_float), ("freq", C.c_float), ("is_top", C.c_int32)]


# every symbol include/peasoup_b200.h declares (checked by tests/test_abi.py)
SYMBOLS = [
    "ps_last_error", "ps_version", "ps_resampleII", "ps_resample", "ps_resample_index_map",
    "ps_form_power_series", "ps_normalise", "ps_harmonic_sum", "ps_peak_workspace_create",
    "ps_peak_workspace_destroy", "ps_find_peaks", "ps_peak_workspace_needed", "ps_median_scrunch5", "ps_linear_stretch",
    "ps_running_median", "ps_running_median_noalloc", "ps_divide_c_by_f", "ps_zap_birdies", "ps_rms", "ps_mean", "ps_fill",
    "ps_convert_u8_f32", "ps_fft_plan_create", "ps_fft_plan_destroy", "ps_fft_r2c", "ps_fft_c2r",
    "ps_search_create", "ps_search_destroy", "ps_search_set_zaplist", "ps_search_whiten",
    "ps_search_accel_batch", "ps_search_reset_peaks", "ps_search_fetch_peaks", "ps_search_peaks_needed",
    "ps_search_reserve_peaks",
    "ps_search_harmonic_peaks", "ps_search_resample_index_map", "ps_search_whitened_series", "ps_search_last_spectrum", "ps_launch_count", "ps_count_transfer",
    "ps_transfer_bytes", "ps_pipeline_search", "ps_pipeline_profile", "ps_pipeline_open",
    "ps_pipeline_run", "ps_pipeline_close", "ps_dedisp_generate_dm_list", "ps_dedisp_max_delay",
    "ps_dedisperse", "ps_dedisperse_on", "ps_search_filterbank", "ps_fold_timeseries", "ps_fold_optimiser_create",
    "ps_fold_optimiser_destroy", "ps_fold_optimise", "ps_fold_results_last", "ps_coincidencer",
    "ps_coincidence_files",
    "ps_accel_plan_generate", "ps_cluster_peaks", "ps_peak_range", "ps_distill", "ps_distill_merged",
]


def build(verbose: bool = False) -> None:
    """Compile libpeasoup_b200.so in-tree with nvcc for sm_100a (see Makefile)."""
    r = subprocess.run(["make", "-C", REPO, "all"], capture_output=True, text=True)
    if verbose or r.returncode:
        print(r.stdout[-4000:], r.stderr[-4000:])
    if r.returncode:
        raise PeasoupError("building libpeasoup_b200.so failed")


def load() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise PeasoupError(f"{LIB_PATH} is missing: run `make` (or __graft_entry__.build()); there is no CPU fallback")
    lib = C.CDLL(LIB_PATH)
    vp, i, u, f, sz, d = C.c_void_p, C.c_int, C.c_uint, C.c_float, C.c_size_t, C.c_double
    lib.ps_last_error.restype = C.c_char_p
    lib.ps_launch_count.restype = C.c_ulonglong
    sig = {
        "ps_resampleII": [vp, vp, sz, f, f, vp], "ps_resample": [vp, vp, sz, f, f, vp],
        "ps_resample_index_map": [vp, sz, f, f, i, vp],
        "ps_form_power_series": [vp, vp, sz, i, vp], "ps_normalise": [vp, f, f, u, vp],
        "ps_harmonic_sum": [vp, C.POINTER(vp), sz, u, vp],
        "ps_peak_workspace_create": [sz, C.POINTER(vp)], "ps_peak_workspace_destroy": [vp],
        "ps_find_peaks": [i, i, vp, f, vp, vp, sz, vp, vp],
        "ps_median_scrunch5": [vp, sz, vp, vp], "ps_linear_stretch": [vp, sz, vp, sz, vp],
        "ps_running_median": [vp, u, d, f, f, vp, vp, vp],
        "ps_running_median_noalloc": [vp, u, d, f, f, vp, vp],
        "ps_divide_c_by_f": [vp, vp, u, vp], "ps_zap_birdies": [vp, vp, vp, f, u, u, vp],
        "ps_rms": [vp, i, i, C.POINTER(f)], "ps_mean": [vp, i, i, C.POINTER(f)], "ps_fill": [vp, vp, f, vp],
        "ps_convert_u8_f32": [vp, vp, u, vp],
        "ps_fft_plan_create": [u, C.POINTER(vp)], "ps_fft_plan_destroy": [vp],
        "ps_fft_r2c": [vp, vp, vp, vp], "ps_fft_c2r": [vp, vp, vp, vp],
        "ps_search_create": [C.POINTER(SearchParams), C.POINTER(vp)], "ps_search_destroy": [vp],
        "ps_search_set_zaplist": [vp, vp, vp, u],
        "ps_search_whiten": [vp, vp, C.POINTER(f), C.POINTER(f), vp],
        "ps_search_accel_batch": [vp, vp, i, i, vp], "ps_search_reset_peaks": [vp, vp],
        "ps_search_fetch_peaks": [vp, vp, sz, vp],
        "ps_search_peaks_needed": [vp], "ps_search_reserve_peaks": [vp, u],
        "ps_search_harmonic_peaks": [vp, vp, sz, i, i, vp],
        "ps_search_whitened_series": [vp], "ps_search_last_spectrum": [vp],
        "ps_search_resample_index_map": [vp, f, vp, C.POINTER(u), vp],
        "ps_pipeline_search": [C.POINTER(PipelineParams), vp, vp, i, i, vp, i, vp, vp, i, i, i, vp, C.c_long,
                               C.POINTER(C.c_ulonglong)],
        "ps_pipeline_open": [C.POINTER(PipelineParams), vp, vp, i, i, i, C.POINTER(vp)],
        "ps_pipeline_run": [vp, vp, vp, i, i, vp, i, vp, C.c_long, C.POINTER(C.c_ulonglong)],
        "ps_pipeline_close": [vp],
        "ps_dedisp_generate_dm_list": [f, f, f, f, f, f, i, f, vp, i],
        "ps_dedisp_max_delay": [vp, i, i, f, f, f],
        "ps_dedisperse": [vp, sz, i, i, f, f, f, vp, i, vp, vp, sz],
        "ps_dedisperse_on": [i, sz, vp, sz, i, i, f, f, f, vp, i, vp, vp, sz],
        "ps_search_filterbank": [C.c_char_p, C.POINTER(FilterbankParams), vp, C.c_long, vp, i, C.POINTER(i), vp],
        "ps_fold_timeseries": [vp, vp, sz, sz, d, d, i, vp],
        "ps_fold_optimiser_create": [u, u, C.POINTER(vp)], "ps_fold_optimiser_destroy": [vp],
        "ps_fold_optimise": [vp, vp, d, f, vp, vp, C.POINTER(FoldResult), vp],
        "ps_fold_results_last": [vp, vp, vp, C.c_long],
        "ps_coincidencer": [vp, vp, i, sz, f, i, vp],
        "ps_coincidence_files": [C.POINTER(C.c_char_p), i, C.POINTER(CoincidenceParams), vp, sz, vp, sz],
        "ps_accel_plan_generate": [f, f, f, f, u, f, f, f, f, vp, i],
        "ps_cluster_peaks": [vp, vp, i, i, vp, vp],
        "ps_peak_range": [i, d, i, f, f, C.POINTER(i), C.POINTER(i), C.POINTER(f)],
        "ps_distill": [i, f, f, i, i, vp, C.c_long, vp, C.c_long],
        "ps_distill_merged": [vp, C.c_long, f, i, vp, C.c_long],
    }
    for name, args in sig.items():
        getattr(lib, name).argtypes = args
    lib.ps_count_transfer.argtypes = [C.c_ulonglong, C.c_ulonglong]
    lib.ps_count_transfer.restype = None
    lib.ps_transfer_bytes.argtypes = [C.POINTER(C.c_ulonglong), C.POINTER(C.c_ulonglong)]
    lib.ps_transfer_bytes.restype = None
    lib.ps_pipeline_profile.argtypes = [C.POINTER(C.c_double)]
    lib.ps_pipeline_profile.restype = None
    lib.ps_search_fetch_peaks.restype = C.c_long
    lib.ps_search_peaks_needed.restype = C.c_uint
    lib.ps_pipeline_search.restype = C.c_long
    lib.ps_pipeline_run.restype = C.c_long
    lib.ps_search_filterbank.restype = C.c_long
    lib.ps_dedisp_max_delay.restype = C.c_long
    lib.ps_pipeline_close.restype = None
    lib.ps_distill.restype = C.c_long
    lib.ps_distill_merged.restype = C.c_long
    lib.ps_fold_results_last.restype = C.c_long
    lib.ps_coincidence_files.restype = C.c_long
    lib.ps_fold_optimiser_destroy.restype = None
    lib.ps_search_whitened_series.restype = vp
    lib.ps_search_last_spectrum.restype = vp
    lib.ps_peak_workspace_destroy.restype = None
    lib.ps_peak_workspace_needed.argtypes = [vp]
    lib.ps_peak_workspace_needed.restype = C.c_size_t
    lib.ps_fft_plan_destroy.restype = None
    lib.ps_search_destroy.restype = None
    _lib = lib
    return lib


def check(rc: int) -> int:
    """Raise PeasoupError (with ps_last_error) on a negative status; pass counts through."""
    if rc < 0:
        raise PeasoupError(load().ps_last_error().decode(errors="replace"))
    return rc


def ptr(t) -> int:
    """Device/host pointer of a torch tensor or numpy array as an int."""
    if hasattr(t, "data_ptr"):
        return t.data_ptr()
    return t.ctypes.data


from .host import (FFTPlan, Pipeline, Search, accel_plan, cluster_peaks, distill, peak_range,  # noqa: E402
                   pipeline_search)
