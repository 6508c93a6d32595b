#!/bin/bash
# SASS listings of the hot kernels (profiles/sass/<kernel>.sass.gz) + a mnemonic census proving the tcgen05 / vector-atomic paths.
set -e
cd "$(dirname "$0")/.."
B=nerf_sys_b200/csrc/build
out=profiles/sass
mkdir -p $out
dump() {  # object, mangled-name regex, label
  fn=$(cuobjdump -elf $B/$1 2>/dev/null | grep -oE "_ZN3acn[A-Za-z0-9_]*$2[A-Za-z0-9_]*" | sort -u | head -1)
  [ -z "$fn" ] && { echo "no function matching $2 in $1"; return; }
  cuobjdump -sass -fun "$fn" $B/$1 | grep -vE "^\s*$" > /tmp/$3.sass
  gzip -9 -c /tmp/$3.sass > $out/$3.sass.gz
  printf "%-28s %6d instructions: " $3 $(grep -cE "^\s+/\*[0-9a-f]{4,6}\*/" /tmp/$3.sass)
  grep -oE "\b(UTCHMMA|UTCBAR|LDTM|STTM|UTCATOMSWS|SYNCS|ATOM|RED|REDG|LDG|LDGSTS|STS|LDS|BAR|FENCE|SHFL|MUFU|HMMA|FFMA)(\.[A-Z0-9_x]+)*" /tmp/$3.sass \
    | sed -E 's/^(UTCHMMA|UTCBAR|LDTM\.x[0-9]+|LDTM|UTCATOMSWS|SYNCS|ATOM\.E\.ADD\.F32x[24]|ATOM|REDG|RED|LDG\.E\.(64|128)|LDG|STS\.128|STS|LDS|BAR|FENCE|SHFL|MUFU|HMMA|FFMA).*/\1/' \
    | sort | uniq -c | sort -rn | awk '{printf "%s x%s  ", $2, $1}'
  echo
}
{
dump ngp_tc.o 16k_ngp_mlp_fwd_tc k_ngp_mlp_fwd_tc
dump ngp_tc.o "16k_ngp_mlp_bwd_tcILi4" k_ngp_mlp_bwd_tc_512
dump ngp_tc.o "16k_ngp_mlp_bwd_tcILi2" k_ngp_mlp_bwd_tc_256
dump dense_tc.o "10k_dense_tcILi64" k_dense_tc_64
dump dense_tc.o 10k_wgrad_tc k_wgrad_tc
dump ngp.o 14k_ngp_hash_fwd k_ngp_hash_fwd
dump ngp.o 14k_ngp_hash_bwd k_ngp_hash_bwd
dump render.o 15k_composite_fwd k_composite_fwd
dump render.o 15k_composite_bwd k_composite_bwd
dump routing.o "13k_route_writeILi8ELb0" k_route_write
dump routing.o "13k_route_countILi8ELb0" k_route_count
dump a2a.o 15k_a2a_push_rows k_a2a_push_rows
dump a2a.o 18k_a2a_pull_combine k_a2a_pull_combine
dump a2a.o 15k_a2a_push_grad k_a2a_push_grad
dump optim.o 12k_adam_multi k_adam_multi
} | tee $out/CENSUS.txt
