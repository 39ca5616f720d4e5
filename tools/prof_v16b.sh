set -x
O=gpurun_out
CMD="python tools/bench_jpeg.py --streams 32768 --reps 1"
ncu --set full --clock-control none --import-source on -k regex:melf_jpeg_huffman -s 1 -c 1 -o $O/v16_jpeg_huffman_sat -f $CMD > $O/ncu_sat.log 2>&1
tail -3 $O/ncu_sat.log
