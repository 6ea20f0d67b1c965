#!/bin/bash
# SASS mnemonic counts per kernel family of the built library: bash tools/sass_summary.sh > profiles/r02_sass_summary.txt
lib=s2p_b200/lib/libs2p_sm100.so
echo "# cuobjdump -sass $lib (sm_100a), mnemonic counts per kernel family; built by make -C s2p_b200/csrc"
echo "# UTCHMMA = tcgen05.mma, UTCBAR = tcgen05.commit, LDTM = tcgen05.ld, UTMALDG = cp.async.bulk.tensor (TMA load), UBLKCP = cp.async.bulk, SYNCS = mbarrier"
cuobjdump -sass $lib 2>/dev/null | awk '
/Function : / { fn=$3; fam=fn;
  if (fn ~ /conv_igemm_kernel/) fam="conv_igemm_kernel"; else if (fn ~ /conv_wgrad_pair_kernel/) fam="conv_wgrad_pair_kernel";
  else if (fn ~ /conv_wgrad_kernel/) fam="conv_wgrad_kernel"; else if (fn ~ /augment_u8_staged/) fam="augment_u8_staged_kernel";
  else if (fn ~ /mat_norm_kernel/) fam="mat_norm_kernel"; else fam="(other)"; inst[fam]++ ; next }
{ for (i=1;i<=NF;i++) { t=$i; sub(/^@!?U?P[0-9T]+/,"",t);
    if (t ~ /^UTCHMMA\.2CTA/) c[fam,"UTCHMMA.2CTA"]++; else if (t ~ /^UTCHMMA/) c[fam,"UTCHMMA"]++;
    else if (t ~ /^UTCBAR/) c[fam,"UTCBAR"]++; else if (t ~ /^LDTM/) c[fam,"LDTM"]++;
    else if (t ~ /^UTMALDG/) c[fam,"UTMALDG"]++; else if (t ~ /^UBLKCP/) c[fam,"UBLKCP"]++;
    else if (t ~ /^SYNCS/) c[fam,"SYNCS"]++; else if (t ~ /^HMMA|^WGMMA/) c[fam,"HMMA/WGMMA"]++; } }
END { n=split("conv_igemm_kernel conv_wgrad_kernel conv_wgrad_pair_kernel augment_u8_staged_kernel mat_norm_kernel (other)", f, " ");
  m=split("UTCHMMA UTCHMMA.2CTA UTCBAR LDTM UTMALDG UBLKCP SYNCS HMMA/WGMMA", k, " ");
  printf "%-28s %6s", "kernel family", "kernels"; for (j=1;j<=m;j++) printf " %13s", k[j]; printf "\n";
  for (i=1;i<=n;i++) { printf "%-28s %6d", f[i], inst[f[i]]; for (j=1;j<=m;j++) printf " %13d", c[f[i],k[j]]; printf "\n" } }'
