for lib in daala_b200/libdaala_tdlt_cuda.so daala_b200/variants/lib_aligned.so; do
  echo "== $lib"
  DAALA_TDLT_LIB=$PWD/$lib timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_op_write.sum,lts__t_sectors_op_read.sum,gpu__time_duration.sum --clock-control none -k regex:tdlt_inv_region_kernel -s 4 -c 1 --csv python tools/ablate.py --one 64 quad 1 2>/dev/null | grep -E "dram__|lts__|gpu__time" | awk -F, '{print $(NF-2), $(NF-1), $NF}'
done
