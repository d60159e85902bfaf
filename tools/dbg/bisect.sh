for k in "split_and_sb" "dct_block_sweep" "lap_filter_sweep" "dering_block" "resample_luma_block" "selftest or check_bias or ieee" "frames_match" "pixel" "dering_frames or smooth" "coding_order" "full_size or 8k" "cfl"; do
  r=$(timeout 600 python -m pytest tests -m gpu -x -q -k "($k) or (dropin and 30-4-4)" 2>&1 | tail -1)
  echo "$k => $r"
done
