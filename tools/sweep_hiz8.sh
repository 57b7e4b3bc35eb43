# development aid (next experiment, not yet measured): the warp-cooperative second pyramid test of k_setup
# (-DKHIZ8=N: boxes of up to N 8-px tiles).  Builds the variants, checks them against the oracle, times them.
#   gpurun -- 'bash tools/sweep_hiz8.sh > gpurun_out/sweep_hiz8.txt 2>&1'
for n in 64 256 1024; do python -m dirt_b200.build -DKHIZ8=$n --out=libvar_hiz8_$n.so > /dev/null || exit 1; done
for lib in libdirt_b200 libvar_hiz8_64 libvar_hiz8_256 libvar_hiz8_1024; do
  export DIRT_B200_LIB=$PWD/dirt_b200/$lib.so
  echo "$lib: $(python -m pytest tests/test_fuzz_gpu.py tests/test_parity_gpu.py tests/test_full_size_gpu.py -m gpu -x -q -k 'ordered or sphere or full_size or plans' 2>&1 | tail -1)"
  echo "$lib: $(python tools/stages.py headline c4 2>&1 | grep -E 'headline|c4' | sed -e 's/clear=.*depth_order=/depth_order=/' -e 's/bin_count.*resolve=/resolve=/' -e 's/grad_clear.*//' | tr '\n' ' ')"
done
