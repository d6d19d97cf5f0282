#!/bin/bash
# Usage: tools/go_vectors/run.sh /path/to/reference   (needs a Go toolchain; the reference tree is only read)
# Builds a scratch module around the UNMODIFIED dpf/*.go of the reference (it has no go.mod of its own) and writes
# tests/golden/go_vectors.json.  Then: python -m pytest tests/test_oracle_golden.py -k go_vectors
set -euo pipefail
REF=$(cd "$1" && pwd)
HERE=$(cd "$(dirname "$0")" && pwd)
ROOT=$(cd "$HERE/../.." && pwd)
W=$(mktemp -d)
mkdir -p "$W/ref/dpf" "$W/gen"
for f in "$REF"/dpf/*.go; do case "$f" in *_test.go) ;; *) cp "$f" "$W/ref/dpf/";; esac; done
printf 'module github.com/sachaservan/pir\n\ngo 1.18\n' > "$W/ref/go.mod"
cp "$HERE/main.go" "$W/gen/"
printf 'module go_vectors\n\ngo 1.18\n\nrequire github.com/sachaservan/pir v0.0.0\n\nreplace github.com/sachaservan/pir => ../ref\n' > "$W/gen/go.mod"
(cd "$W/gen" && go run . "$ROOT/tests/golden/dpf_pir_vectors.json") > "$ROOT/tests/golden/go_vectors.json"
echo "wrote tests/golden/go_vectors.json"
