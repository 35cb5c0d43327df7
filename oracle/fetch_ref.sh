#!/bin/bash
# Installs the UNMODIFIED reference package (geduardo/SPARC, `wedm`) into oracle/_ref/ so that bench.py can
# time the reference's own `WireEDMEnv.step` on the GPU box's host cores (BASELINE.json: "vs ref CPU
# env.step"; /root/reference does not exist there).  oracle/_ref/ is git-ignored (no reference source ever
# enters the history) but not gpurun-ignored, so it travels with the snapshot like a built .so.
# Test / measurement infrastructure only: nothing under sparc_b200/ imports it.
set -u
here="$(cd "$(dirname "$0")" && pwd)"
ref="${1:-/root/reference}"
if [ ! -d "$ref/src/wedm" ]; then echo "fetch_ref: $ref/src/wedm not found (nothing to do on the GPU box)"; exit 0; fi
rm -rf "$here/_ref"; mkdir -p "$here/_ref"
tmp="$(mktemp -d)"; cp -r "$ref" "$tmp/ref"          # the build writes into the source tree; /root/reference is read-only
if python -m pip install --quiet --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
        --target "$here/_ref" "$tmp/ref" > "$here/_ref/install.log" 2>&1 && [ -d "$here/_ref/wedm" ]; then
    echo "fetch_ref: pip-installed the reference into oracle/_ref"
else
    # the package is pure Python: its installed form IS its source directory
    cp -r "$ref/src/wedm" "$here/_ref/wedm"
    echo "fetch_ref: pip install failed (see oracle/_ref/install.log); copied the pure-Python package instead"
fi
rm -rf "$tmp"
ls "$here/_ref"
