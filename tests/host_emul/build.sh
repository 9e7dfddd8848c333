#!/bin/sh
# Builds the TEST-ONLY host emulation of the tick kernels (see emul.cc). No FMA contraction, like the x86-64 reference.
set -e
cd "$(dirname "$0")"
/usr/bin/g++ -O2 -std=c++17 -fPIC -shared -ffp-contract=off -w -o libhrg_emul.so emul.cc -lm
