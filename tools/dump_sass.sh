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
dump ngp_tc.o "16k_ngp_mlp_fwd_tcILi7ELi72ELb0" k_ngp_mlp_fwd_tc
dump ngp_tc.o "16k_ngp_mlp_fwd_tcILi5ELi96ELb1" k_ngp_fused_fwd_tc
dump ngp_tc.o "16k_ngp_mlp_bwd_tcILi4" k_ngp_mlp_bwd_tc_512
dump ngp_tc.o "16k_ngp_mlp_bwd_tcILi2" k_ngp_mlp_bwd_tc_256
dump ngp_tc.o 17k_tc_wgrad_reduce k_tc_wgrad_reduce
dump ngp_tc.o 8k_absmax k_absmax
dump dense_tc.o "10k_dense_tcILi64" k_dense_tc_64
dump dense_tc.o 10k_wgrad_tc k_wgrad_tc
dump dense_tc.o 6k_pe_h k_pe_h
dump dense_tc.o 14k_pe_final_fwd k_pe_final_fwd
dump dense_tc.o 14k_pe_final_bwd k_pe_final_bwd
dump ngp.o "14k_ngp_hash_fwdILi4" k_ngp_hash_fwd_4
dump ngp.o "14k_ngp_hash_fwdILi8" k_ngp_hash_fwd_8
dump ngp.o 14k_ngp_hash_bwd k_ngp_hash_bwd
dump render.o "15k_composite_fwdEPK" k_composite_fwd
dump render.o 18k_composite_fwd_sm k_composite_fwd_sm
dump render.o "15k_composite_bwdILi32" k_composite_bwd
dump render.o 8k_bg_fwd k_bg_fwd
dump render.o 14k_bg_bwd_fused k_bg_bwd_fused
dump render.o 13k_mse_partial k_mse_partial
dump rays.o 8k_raygen k_raygen
dump rays.o 9k_strat_t k_strat_t
dump routing.o "13k_route_writeILi8EE" k_route_write
dump routing.o "13k_route_countILi8ELb0" k_route_count
dump a2a.o 15k_a2a_push_rows k_a2a_push_rows
dump a2a.o 18k_a2a_pull_combine k_a2a_pull_combine
dump a2a.o 15k_a2a_push_grad k_a2a_push_grad
dump a2a.o 13k_p2p_barrier k_p2p_barrier
dump a2a.o 10k_p2p_plan k_p2p_plan
dump optim.o 12k_adam_multi k_adam_multi
dump optim.o 13k_sumsq_multi k_sumsq_multi
dump occ.o 11k_occ_march k_occ_march
dump occ.o 11k_seg_union k_seg_union
dump occ.o 12k_packed_fwd k_packed_fwd
dump occ.o 12k_packed_bwd k_packed_bwd
dump episode.o 15k_episode_route k_episode_route
dump voronoi.o 9k_voronoi k_voronoi
} | tee $out/CENSUS.txt
