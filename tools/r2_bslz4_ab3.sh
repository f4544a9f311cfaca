export FFS_DEV_KNOBS=1
O=gpurun_out/bslz4ab3; mkdir -p $O
V=$PWD/fast-feedback-service_b200/ffs_b200/variants
( FFS_B200_LIB=$V/libffs_b200_pers.so timeout 300 python -m pytest tests/test_gpu_bslz4.py tests/test_gpu_cli.py -x -q -m gpu 2>&1 | tail -2 )
( timeout 300 python -m pytest tests/test_gpu_bslz4.py -x -q -m gpu 2>&1 | tail -1 )
for n in cur pers cur pers; do
  lib=$V/libffs_b200_$n.so; [ "$n" = cur ] && lib=$PWD/fast-feedback-service_b200/ffs_b200/libffs_b200.so
  for bpp in 2 4; do
    FFS_B200_LIB=$lib BPP=$bpp timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:bslz4_decode -s 2 -c 5 --csv --log-file $O/ncu_${n}_$bpp.csv python tools/bench_bslz4.py > $O/ncu_${n}_$bpp.log 2>&1
    echo "$n bpp=$bpp: $(grep -E "bslz4_decode" $O/ncu_${n}_$bpp.csv | awk -F'","' '{print $NF}' | tr -d '"' | paste -sd' ')"
  done
done
