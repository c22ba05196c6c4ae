#!/bin/sh
# TEST INFRASTRUCTURE ONLY: builds the product sources against the host emulation of CUDA.
set -e
cd "$(dirname "$0")/../.."
mkdir -p tests/_emu
g++ -std=c++17 -O2 -g -fPIC -shared -Wl,-Bsymbolic -DNP_EMU -Itests/emu -Iincflo_b200/csrc -Wall -Wno-unused-function \
    -Wno-unknown-pragmas -o tests/_emu/libnodalproj_emu.so tests/emu/nodalproj_emu.cpp
