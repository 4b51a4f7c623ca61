mkdir -p gpurun_out/s14
A=""
for g in tn2x64/kb8x16/w8 tn4x64/kb8x16/w8 tn4x64/kb4x16/w8 tn8x64/kb4x16/w8 tn4x64/kb8x16/w16 tn2x64/kb8x16/w16 tn2x64/kb16x16/w8; do
  P=__half/nt2/$g
  A="$A --combo qkv=$P,o=$P,gate_up=$P,down=$P"
done
python tools/chain.py --m 16 --combo "" $A > gpurun_out/s14/chain_m16_v3.txt 2>&1
B=""
for c in "qkv=st/tn128/kb128/w8/S6,o=st/tn128/kb128/w8/S8,gate_up=st/tn256/kb128/w8/S2,down=st/tn128/kb128/w8/S8" "qkv=st/tn128/kb128/w4/S4,o=st/tn128/kb128/w4/S8,gate_up=st/tn128/kb128/w4/S1,down=st/tn128/kb128/w4/S8" "qkv=st/tn64/kb128/w4/S5,o=st/tn64/kb128/w4/S6,gate_up=st/tn256/kb128/w8/s3/S3,down=st/tn64/kb128/w4/S6"; do B="$B --combo $c"; done
python tools/chain.py --m 16 $B > gpurun_out/s14/chain_m16_v2.txt 2>&1
cut -c1-130 gpurun_out/s14/chain_m16_v3.txt gpurun_out/s14/chain_m16_v2.txt
A=""
for g in tn2x64/kb8x16/w8 tn4x64/kb8x16/w8 tn8x64/kb4x16/w8 tn4x64/kb8x16/w16; do
  P=__half/nt4/$g
  A="$A --combo qkv=$P,o=$P,gate_up=$P,down=$P"
done
python tools/chain.py --m 32 --combo "" $A --combo "qkv=1008,o=1008,gate_up=1005,down=1008" --combo "qkv=1007,o=1008,gate_up=1005,down=1008" > gpurun_out/s14/chain_m32.txt 2>&1
cut -c1-130 gpurun_out/s14/chain_m32.txt
python tools/timeline.py --m 16 --layers 2 > gpurun_out/s14/tl_m16.txt 2>&1; tail -n 6 gpurun_out/s14/tl_m16.txt
