"""The C-ABI library loads without a GPU and exports every symbol include/peasoup_b200.h declares."""
import os
import re

import peasoup_b200
from conftest import REPO


def declared_symbols():
    txt = open(os.path.join(REPO, "include", "peasoup_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(ps_[a-z0-9_A-Z]+)\s*\(", txt)))


def test_header_and_binding_agree():
    assert declared_symbols() == sorted(peasoup_b200.SYMBOLS)


def test_library_exports_every_declared_symbol(lib):
    for name in declared_symbols():
        assert hasattr(lib, name), name


def test_version_and_error_string(lib):
    assert lib.ps_version() >= 100
    assert isinstance(lib.ps_last_error(), bytes)


def test_every_declaration_cites_the_reference():
    txt = open(os.path.join(REPO, "include", "peasoup_b200.h")).read()
    for cite in ["kernels.h:126", "kernels.h:118", "kernels.h:111", "kernels.h:144", "kernels.h:104", "kernels.h:134",
                 "kernels.h:252", "kernels.h:262", "kernels.h:268", "kernels.h:279", "kernels.h:318", "ffter.hpp:46-78",
                 "pipeline_multi.cu:100-252", "utils.hpp:168-192", "peakfinder.hpp:27-56", "distiller.hpp"]:
        assert cite in txt, cite
