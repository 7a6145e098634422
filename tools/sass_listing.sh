#!/bin/bash
# Dumps the SASS of the tensor-core / TMA kernels (objects built by __graft_entry__.build()); profiles/r2/sass_tensor_tma_kernels.txt
# holds the mnemonic counts extracted from these dumps.
for f in corr_tc2 lookup_tc corr_tcgen05 lookup; do cuobjdump -sass cardiac-segmentation-optical-flow_b200/build/$f.o > /tmp/$f.sass; done
grep -cE "UTCHMMA|LDTM|UTMALDG|UTMASTG|UTCBAR|LDGSTS" /tmp/corr_tc2.sass /tmp/lookup_tc.sass /tmp/corr_tcgen05.sass /tmp/lookup.sass
