#!/bin/bash
# Stage the UNMODIFIED reference (pure Python) under oracle/_ref/ so that it travels to the GPU box with the
# gpurun snapshot (oracle/_ref/ is git-ignored, not gpurun-ignored).  TEST / BASELINE INFRASTRUCTURE ONLY:
# used by tests/ (drop-in checks), by `bench.py --impl reference` and by bench.py's `cpu_baseline` /
# `gpu_reference` legs.  Nothing under calodiffusion_b200/ imports it.
#
#   bash oracle/build_ref.sh            # needs /root/reference (build container); no-op when it is absent
#
# What is copied (verbatim, no edits): the `calodiffusion` package (models, samplers, utils, configs), the
# CaloChallenge binning XMLs + helper modules it reads, and HGCalShowers (HGCalGeo.py + geoms/geom_william.pkl,
# the geometry BASELINE config 4 names).  Byte-compiled caches and notebooks are left out.
set -e
SRC=${1:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
DST="$HERE/_ref"
if [ ! -d "$SRC/calodiffusion" ]; then
  echo "build_ref: $SRC/calodiffusion not found; keeping existing $DST (if any)"
  exit 0
fi
rm -rf "$DST"
mkdir -p "$DST"
for d in calodiffusion CaloChallenge HGCalShowers; do
  (cd "$SRC" && find "$d" -type f ! -name '*.pyc' ! -name '*.ipynb' ! -name '*.jpg' ! -name '*.png' ! -path '*/__pycache__/*' -print0) |
    (cd "$SRC" && xargs -0 cp --parents -t "$DST")
done
# provenance: file list + sha256 of every staged file, so a reader can check nothing was edited
(cd "$DST" && find . -type f ! -name MANIFEST.sha256 -print0 | sort -z | xargs -0 sha256sum) > "$DST/MANIFEST.sha256"
echo "build_ref: staged $(wc -l < "$DST/MANIFEST.sha256") files from $SRC into $DST ($(du -sh "$DST" | cut -f1))"
