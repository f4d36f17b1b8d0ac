#!/bin/sh
# Copies the reference's Python package, untouched, into the git-ignored oracle/_ref/ so that the
# CPU baseline of bench.py can run the reference itself on the GPU box (which has no
# /root/reference). Build container only. Same recipe as __graft_entry__.build().
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
SRC="${1:-/root/reference}"
[ -d "$SRC/collie" ] || { echo "no reference at $SRC"; exit 0; }
rm -rf "$HERE/_ref"
mkdir -p "$HERE/_ref"
cp -r "$SRC/collie" "$HERE/_ref/collie"
find "$HERE/_ref" -name '__pycache__' -type d -prune -exec rm -rf {} +
echo "reference copied to $HERE/_ref/collie"
