#!/bin/bash
# TEST INFRASTRUCTURE ONLY -- builds the *reference* overlapInCore (and the three
# reference tools the parity harness needs) straight from the sources where they lie
# under /root/reference/src, into oracle/_ref/ (git-ignored).  Nothing is copied into
# the repo; the reference's own build system (src/Makefile, boilermake) is NOT run:
# this is a plain g++ loop over the few translation units the path needs.
#
# Two portability shims are generated under oracle/_ref/shim/ (build outputs):
#   - sys/sysctl.h        empty; AS_UTL/AS_UTL_alloc.C:40 includes a header that modern
#                         glibc no longer ships (the sysconf branch is what compiles)
#   - canu_version.H      the header src/canu_version_update.pl would generate
# and -fpermissive downgrades the ill-formed, never-instantiated accessors in
# AS_UTL/stddev.H:341,346,425,430 to warnings.  None of this touches the hot path.
#
# Flags follow src/Makefile:387-402 (no -march=native, no -ffast-math).
set -e
REF=${REF:-/root/reference/src}
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
OBJ="$OUT/obj"
CXX=${ORACLE_CXX:-/usr/bin/g++}

if [ ! -d "$REF" ]; then
  echo "build_ref.sh: $REF not present (GPU box?) -- using prebuilt $OUT" >&2
  exit 0
fi
if [ -x "$OUT/overlapInCore" ] && [ -x "$OUT/gatekeeperCreate" ] && [ -x "$OUT/overlapConvert" ] \
   && [ -x "$OUT/overlapInCorePartition" ] && [ -x "$OUT/meryl" ] && [ -x "$OUT/ovStoreBuild" ] && [ -x "$OUT/ovStoreDump" ] \
   && [ "$1" != "-f" ]; then
  exit 0
fi

mkdir -p "$OBJ" "$OUT/shim/sys"
: > "$OUT/shim/sys/sysctl.h"
cat > "$OUT/shim/canu_version.H" <<'EOF'
#define CANU_VERSION_LABEL     "snapshot"
#define CANU_VERSION_MAJOR     "1"
#define CANU_VERSION_MINOR     "6"
#define CANU_VERSION_COMMITS   "0"
#define CANU_VERSION_REVISION  "0"
#define CANU_VERSION_HASH      ""
#define CANU_VERSION           "Canu 1.6\n"
EOF

CXXFLAGS="-D_GLIBCXX_PARALLEL -pthread -fopenmp -fPIC -w -fpermissive \
 -O3 -funroll-loops -fexpensive-optimizations -finline-functions -fomit-frame-pointer"
INC="-I$OUT/shim -I$REF -I$REF/AS_UTL -I$REF/stores -I$REF/stores/libsnappy \
 -I$REF/overlapInCore -I$REF/overlapInCore/liboverlap -I$REF/meryl/libleaff"

LIB_SRCS="
AS_global.C
AS_UTL/AS_UTL_decodeRange.C AS_UTL/AS_UTL_fasta.C AS_UTL/AS_UTL_fileIO.C
AS_UTL/AS_UTL_reverseComplement.C AS_UTL/AS_UTL_stackTrace.C AS_UTL/AS_UTL_alloc.C
AS_UTL/bitEncodings.C AS_UTL/bitPackedFile.C AS_UTL/bitPackedArray.C AS_UTL/dnaAlphabets.C
AS_UTL/hexDump.C AS_UTL/md5.C AS_UTL/mt19937ar.C AS_UTL/readBuffer.C AS_UTL/speedCounter.C
AS_UTL/sweatShop.C AS_UTL/timeAndSize.C AS_UTL/kMer.C
stores/gkStore.C stores/gkStoreEncode.C
stores/ovOverlap.C stores/ovStore.C stores/ovStoreWriter.C stores/ovStoreFilter.C
stores/ovStoreFile.C stores/ovStoreHistogram.C
stores/libsnappy/snappy-sinksource.cc stores/libsnappy/snappy-stubs-internal.cc stores/libsnappy/snappy.cc
overlapInCore/liboverlap/Binomial_Bound.C overlapInCore/liboverlap/Display_Alignment.C
overlapInCore/liboverlap/prefixEditDistance.C overlapInCore/liboverlap/prefixEditDistance-allocateMoreSpace.C
overlapInCore/liboverlap/prefixEditDistance-extend.C overlapInCore/liboverlap/prefixEditDistance-forward.C
overlapInCore/liboverlap/prefixEditDistance-reverse.C
"
OIC_SRCS="
overlapInCore/overlapInCore.C overlapInCore/overlapInCore-Build_Hash_Index.C
overlapInCore/overlapInCore-Find_Overlaps.C overlapInCore/overlapInCore-Output.C
overlapInCore/overlapInCore-Process_Overlaps.C overlapInCore/overlapInCore-Process_String_Overlaps.C
"

compile() {  # src -> obj path (echoed)
  local s="$1"; local o="$OBJ/$(echo "$s" | tr '/' '_').o"
  if [ ! -f "$o" ] || [ "$REF/$s" -nt "$o" ]; then
    $CXX $CXXFLAGS $INC -c "$REF/$s" -o "$o" || exit 1
  fi
  echo "$o"
}

# meryl (the k-mer counter upstream of overlapInCore: `meryl -B -C ...` then `meryl -Dt -n T` writes the -k skip list,
# Meryl.pm:434-450,634) -- the oracle of the frequent-k-mer row
MERYL_SRCS="
meryl/meryl-args.C meryl/meryl-binaryOp.C meryl/meryl-build.C meryl/meryl-dump.C meryl/meryl-estimate.C meryl/meryl-merge.C
meryl/meryl-unaryOp.C meryl/meryl.C meryl/libmeryl.C
meryl/libleaff/fastaFile.C meryl/libleaff/fastaStdin.C meryl/libleaff/fastqFile.C meryl/libleaff/fastqStdin.C
meryl/libleaff/gkStoreFile.C meryl/libleaff/merStream.C meryl/libleaff/seqCache.C meryl/libleaff/seqFactory.C
meryl/libleaff/seqStore.C meryl/libleaff/seqStream.C meryl/libleaff/sffFile.C
"
INC="$INC -I$REF/meryl"

pids=""; LIB_OBJS=""
# ovStoreBuild / ovStoreDump: the step downstream of overlapInCore -- oracle of the overlap-store writer
for s in $LIB_SRCS $OIC_SRCS $MERYL_SRCS stores/gatekeeperCreate.C overlapInCore/overlapConvert.C overlapInCore/overlapInCorePartition.C \
         stores/ovStoreBuild.C stores/ovStoreDump.C; do
  ( compile "$s" >/dev/null ) &
  pids="$pids $!"
  # bound parallelism to 8
  if [ $(jobs -r | wc -l) -ge 8 ]; then wait -n; fi
done
wait
for s in $LIB_SRCS; do LIB_OBJS="$LIB_OBJS $(compile $s)"; done
OIC_OBJS=""; for s in $OIC_SRCS; do OIC_OBJS="$OIC_OBJS $(compile $s)"; done

LD="-pthread -fopenmp -lm"
$CXX -o "$OUT/overlapInCore"          $OIC_OBJS $LIB_OBJS $LD
$CXX -o "$OUT/gatekeeperCreate"       "$(compile stores/gatekeeperCreate.C)" $LIB_OBJS $LD
$CXX -o "$OUT/overlapConvert"         "$(compile overlapInCore/overlapConvert.C)" $LIB_OBJS $LD
$CXX -o "$OUT/overlapInCorePartition" "$(compile overlapInCore/overlapInCorePartition.C)" $LIB_OBJS $LD
$CXX -o "$OUT/ovStoreBuild"           "$(compile stores/ovStoreBuild.C)" $LIB_OBJS $LD
$CXX -o "$OUT/ovStoreDump"            "$(compile stores/ovStoreDump.C)" $LIB_OBJS $LD
MERYL_OBJS=""; for s in $MERYL_SRCS; do MERYL_OBJS="$MERYL_OBJS $(compile $s)"; done
$CXX -o "$OUT/meryl"                  $MERYL_OBJS $LIB_OBJS $LD
echo "build_ref.sh: built reference tools in $OUT" >&2
