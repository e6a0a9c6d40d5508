#!/bin/bash
# Compile ONE board size of the library and print registers / stack of its kernels (no GPU needed).
# usage: scripts/dev_regs.sh <S> [kernel-substring] [extra nvcc flags...]
S=${1:-8}; PAT=${2:-k_}; shift 2 2>/dev/null
OUT=${TMPDIR:-/tmp}/oth_dev_$S.o
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -DOTH_S=$S "$@" -c "$(dirname "$0")/../gymothelloenv_b200/csrc/oth_size.cu" -o $OUT || exit 1
cuobjdump -res-usage $OUT | grep -A1 "Function" | paste - - | grep "$PAT" | sed 's/SHARED.*//; s/Function //' | awk '{print $1, $2, $3}'
