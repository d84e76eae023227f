#!/bin/bash
# SASS of the four kernels of the acceleration-trial chain at the bench size (2^23 samples), without the
# encoding columns, into profiles/<prefix>_sass_<kernel>.txt.  Needs build/search.o (make).
# Check e.g.: grep -c UBLKCP profiles/r02_sass_fft_pass2.txt (TMA bulk copies), SYNCS (mbarrier).
set -e
cd "$(dirname "$0")/.."
prefix=${1:-r02}
dump() {  # name, mangled function
  cuobjdump -sass -fun "$2" build/search.o | sed -E 's@ */\* 0x[0-9a-f]+ \*/$@@' | grep -vE '^\s*/\* 0x[0-9a-f]+ \*/$' > profiles/${prefix}_sass_$1.txt
  echo "$1: $(grep -cE '^\s+/\*[0-9a-f]{4}\*/' profiles/${prefix}_sass_$1.txt) instructions"
}
dump fft_pass1 _ZN2ps16fft_pass1_kernelILi8ELb0ENS_13LoadResampledEEEvT1_PN5psfft2cfENS_9FftTablesEiiii
dump fft_pass2 _ZN2ps16fft_pass2_kernelILi8ELb0ENS_12StoreNaturalEEEvPKN5psfft2cfET1_NS_9FftTablesEiiii
dump spectrum _ZN2ps15spectrum_kernelEPK6float2PfjmPKffPKN5psfft2cfES9_S0_
dump harmonic_peaks_rows _ZN2ps26harmonic_peaks_rows_kernelILi4ELi1024ELi8ELi397EEEvPKfimNS_11LevelRangesENS_11LevelScalesEfiNS_8PeakSinkEiiPyy
