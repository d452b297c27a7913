#!/bin/bash
# SASS evidence for the kernels of record: one full listing per kernel (profiles/sass/*.sass) and a table of the
# Blackwell-specific mnemonics each contains (profiles/r2_sass_mnemonics.txt). Runs without a GPU.
SO=collective_communication_mpi_b200/lib/libccmpi_b200.so
OUT=profiles/sass
mkdir -p $OUT
declare -A K=(
 [gemm_plain_256_kn]='_ZN4gemm16gemm_kernel_2ctaILi256ELb0ELb1E13__nv_bfloat16NS_11PlainPolicyELi2ELb0EEEvNS_6LaunchE'
 [gemm_sgd_256_dw]='_ZN4gemm16gemm_kernel_2ctaILi256ELb1ELb1E13__nv_bfloat16NS_9SgdPolicyELi2ELb0EEEvNS_6LaunchE'
 [ag_gemm_256_kn]='_ZN4gemm16gemm_kernel_2ctaILi256ELb0ELb1E13__nv_bfloat16NS_9AgKPolicyELi2ELb0EEEvNS_6LaunchE'
 [gemm_rs_256_kn]='_ZN4gemm16gemm_kernel_2ctaILi256ELb0ELb1E13__nv_bfloat16NS_9RsNPolicyELi2ELb0EEEvNS_6LaunchE'
 [gemm_ar_256_kn]='_ZN4gemm16gemm_kernel_2ctaILi256ELb0ELb1E13__nv_bfloat16NS_8ArPolicyELi2ELb0EEEvNS_6LaunchE'
 [gemm_ag_out_256_kn]='_ZN4gemm16gemm_kernel_2ctaILi256ELb0ELb1E13__nv_bfloat16NS_11AgOutPolicyELi2ELb0EEEvNS_6LaunchE'
 [ag_gemm_mxfp8]='_ZN4gemm16gemm_kernel_2ctaILi128ELb0ELb0E13__nv_bfloat16NS_9AgKPolicyELi1ELb1EEEvNS_6LaunchE'
 [gemm_rs_mxfp8]='_ZN4gemm16gemm_kernel_2ctaILi128ELb0ELb0E13__nv_bfloat16NS_9RsNPolicyELi1ELb1EEEvNS_6LaunchE'
 [allreduce_nvls_bf16]='_ZN4coll21allreduce_nvls_kernelI13__nv_bfloat16S1_EEvNS_6LaunchEmmmf'
 [allreduce_sgd_nvls]='_ZN4coll25allreduce_sgd_nvls_kernelENS_6LaunchEmmmfyjj'
 [allreduce_twoshot_tight_bf16_n2]='_ZN4coll30allreduce_twoshot_tight_kernelI13__nv_bfloat16Li0ES1_Li2EEEvNS_6LaunchEmmmmf'
)
: > profiles/r2_sass_mnemonics.txt
printf "%-34s %6s | %s\n" kernel lines "UTCHMMA UTCQMMA UTCCP UTMALDG UBLKCP LDTM multimem(LDGMC/STGMC/REDG.MC) peer-flag(ATOMG/RED .SYS)" >> profiles/r2_sass_mnemonics.txt
for name in "${!K[@]}"; do
  f=$OUT/$name.sass
  cuobjdump -sass -fun "${K[$name]}" $SO 2>/dev/null | python scripts/_sass_compact.py > $f
  n=$(wc -l < $f)
  c() { grep -c "$1" $f; }
  printf "%-34s %6d | %d %d %d %d %d %d %d %d\n" $name $n $(c UTCHMMA) $(c UTCQMMA) $(c UTCCP) $(c UTMALDG) $(c UBLKCP) $(c LDTM) $(grep -c "LDGMC\|STGMC\|\.MC\b\|MULTIMEM" $f) $(grep -c "\.SYS" $f) >> profiles/r2_sass_mnemonics.txt
done
sort -o profiles/r2_sass_mnemonics.txt profiles/r2_sass_mnemonics.txt
cat profiles/r2_sass_mnemonics.txt; du -sh $OUT
