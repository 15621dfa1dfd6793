#!/bin/sh
# TEST INFRASTRUCTURE ONLY -- builds oracle/_ref/cnf2ref from the reference sources WHERE THEY LIE.
#
# The reference translation unit is streamed from /root/reference into g++ (lines 1-6794, i.e. everything
# before the Boost.Spirit parsers and main()), followed by our driver
# oracle/ref_driver.cpp.  The only textual substitutions are inside the #include "settings.h" expansion:
#   DOTOULBAR 1 -> 0   (toulbar2 binary is not present; changes only the inversion stage, not rows a1-a5)
#   XSTDBITSET 1 -> 0  (xstd/bit_set.hpp is not vendored by the reference)
#   INDCOUNT           (1,000,000 thread-private maps per thread are pointless for oracle-sized inputs)
# cnf2ref_selfing (row g1) additionally flips the compile-time knob SELFING to true and un-comments the two extra
# TYPESEXES / TYPEGENS entries settings.h:21-23 keeps in comments for it (256 states, 192 valid).
# Nothing from /root/reference is written into the repository; the outputs are the binaries.
#
# usage: oracle/build_ref.sh [INDCOUNT]    (default 20000)
set -e
REF=${CNF2_REFERENCE_DIR:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
INDCOUNT=${1:-20000}
OUT="$HERE/_ref"
if [ ! -f "$REF/cnF2freq.cpp" ]; then
	echo "build_ref.sh: $REF/cnF2freq.cpp not present; keeping prebuilt $OUT (if any)" >&2
	exit 0
fi
mkdir -p "$OUT"
SELFING=0
settings() {
	sed -e 's/^#define DOTOULBAR 1/#define DOTOULBAR 0/' \
	    -e 's/^#define XSTDBITSET 1/#define XSTDBITSET 0/' \
	    -e "s/^const int INDCOUNT = [0-9]*;/const int INDCOUNT = $INDCOUNT;/" \
	    "$REF/settings.h" |
	if [ "$SELFING" = 1 ]; then
		sed -e 's/^const bool SELFING = false;/const bool SELFING = true;/' \
		    -e 's|{ 0, 0, 1, 1, 0, 1/\*, 0, 0\*//\*, 0\*/ }|{ 0, 0, 1, 1, 0, 1, 0, 0 }|' \
		    -e 's|{ 1, 0, 0, 1, 0, 0/\*, 2, 2\*//\*, 3\*/ }|{ 1, 0, 0, 1, 0, 0, 2, 2 }|'
	else
		cat
	fi
}
build() {
	# $1 = output name, rest = extra flags
	name=$1
	shift
	{
		echo '#line 1 "reference/cnF2freq.cpp"'
		head -n 139 "$REF/cnF2freq.cpp"
		echo '#line 1 "reference/settings.h"'
		settings
		echo '#line 141 "reference/cnF2freq.cpp"'
		sed -n '141,6794p' "$REF/cnF2freq.cpp"
		echo '#line 1 "oracle/ref_driver.cpp"'
		cat "$HERE/ref_driver.cpp"
	} | g++ -x c++ -std=c++20 -fopenmp -DDOEXTERNFORGCC -w -I "$HERE/refshim" "$@" - -o "$OUT/$name"
}
# sanity: line 140 of the reference must be the settings include we splice around
sed -n '140p' "$REF/cnF2freq.cpp" | grep -q '#include "settings.h"' || { echo "reference layout changed" >&2; exit 1; }
# strict IEEE build: the parity oracle (single-threaded use)
build cnf2ref -O2
# demo.sh:6 flags: the timed CPU baseline
build cnf2ref_fast -O2 -ffast-math
# row g1: the SELFING = true instantiation (strict IEEE, parity)
SELFING=1
settings | grep -q '^const bool SELFING = true;' || { echo "settings.h layout changed (SELFING)" >&2; exit 1; }
settings | grep -q '{ 1, 0, 0, 1, 0, 0, 2, 2 }' || { echo "settings.h layout changed (TYPEGENS)" >&2; exit 1; }
build cnf2ref_selfing -O2
SELFING=0
echo "built $OUT/cnf2ref $OUT/cnf2ref_fast $OUT/cnf2ref_selfing (INDCOUNT=$INDCOUNT)"
