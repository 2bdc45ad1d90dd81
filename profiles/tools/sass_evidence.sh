#!/bin/bash
# SASS mnemonics behind the claims of DESIGN.md, from the built library (no GPU needed):
#   profiles/tools/sass_evidence.sh > profiles/r02_sass.md
LIB=${1:-instag_b200/libigs_raster.so}
T=$(mktemp); cuobjdump -sass "$LIB" > "$T"
fn() { awk -v pat="$1" '/Function : /{f=($0 ~ pat)} f' "$T"; }
show() { # kernel-regex, instruction-regex, title
  echo; echo "### $3"; echo; echo '```'
  fn "$1" | grep -E "$2" | sed -E 's/^ *\/\*[0-9a-f]+\*\/ *//; s/ *\/\* 0x[0-9a-f]+ \*\/ *$//' | sed -E 's/R[0-9]+/R#/g; s/UR[0-9]+/UR#/g; s/P[0-9]/P#/g' | sort | uniq -c | sort -rn | head -${4:-6}
  echo '```'
}
echo "# SASS evidence ($(basename $LIB), sm_100a, $(date -u +%F)); register numbers replaced by #, count = occurrences in the kernel"
show "render_backward_kernelILi1" "HMMA" "render_backward_kernel<1>: pixel sums on the tensor cores (mma.sync m16n8k8 TF32, hi/lo split)"
show "render_backward_kernelILi1" "REDG" "render_backward_kernel<1>: 8-byte vector reductions into the accumulator rows (red.global.add.v2.f32)"
show "render_backward_kernelILi1" "MUFU" "render_backward_kernel<1>: exp / reciprocal"
show "render_backward_kernelILi1" "LDGSTS|LDGDEPBAR|DEPBAR" "render_backward_kernel<1>: list ids (4 bytes) and the survivors' colour / normal planes (16 bytes) by cp.async; wait_group 1 = all but the ids"
show "render_kernelILi1" "LDGSTS|LDGDEPBAR|DEPBAR" "render_kernel<1>: colour / normal planes by cp.async (LDGSTS) one chunk ahead"
show "render_kernelILi1" "ACQBULK|PREEXIT|ATOMG" "render_kernel<1>: griddepcontrol.wait (ACQBULK), unit ticket (ATOMG)"
show "preprocess_kernelILb0ELb1" "ACQBULK|ATOMS|REDG|DFMA" "preprocess_kernel<precomp, smem histogram>: dependent launch, shared-memory tile histogram (ATOMS), flush (REDG), ndc2Pix in double (DFMA)"
show "duplicate_kernelILb1ELb1" "ACQBULK|ATOMS|ATOMG|STG" "duplicate_kernel<smem, own scan>: slot reservation (ATOMG), hand-out (ATOMS), 8-byte words (STG.E.64)"
show "tile_sort_kernel" "ATOMS|CREDUX|LDG|STG|BAR" "tile_sort_kernel: one read of the segment (LDG.E.64), CTA min/max (CREDUX), counting / scatter in shared memory (ATOMS)" 8
show "nvls_allreduce" "LDGMC|STG|ST\.|ATOMG|MEMBAR|ACQBULK|MC" "nvls_allreduce_kernel: multimem.ld_reduce (LDGMC...ADD.F32x4), multimem.st, system-scope CAS flags" 10
rm -f "$T"
