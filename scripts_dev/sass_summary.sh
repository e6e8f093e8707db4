#!/bin/bash
# opcode counts of the built library -> stdout (profiles/r2_sass_opcodes.txt)
for op in UTCHMMA UTCBAR LDTM STTM UTMALDG UTMASTG UTMAREDG USETMAXREG "IDP.2A" "VIMNMX.U16x2" LDGSTS SYNCS; do echo "$op $(cuobjdump -sass tmat_b200/libtmat_b200.so | grep -c "$op")"; done
