# development aid: depth pre-load through L1 (.ca) or L2 only (.cg), with and without pre-load in the nearest bucket
for lib in libdirt_b200 libvar_cg libvar_cgpre libvar_capre; do
  for plan in "0.35;1" "0.1,0.35;1" "0.05,0.15,0.35;1" "0.2;1" "0.5;1"; do
    echo "$lib BUCKETS=$plan: $(DIRT_B200_LIB=$PWD/dirt_b200/$lib.so DIRT_B200_BUCKETS=$plan python tools/stages.py headline c4 2>&1 | grep -E 'headline|c4' | sed -e 's/clear=.*depth_order=/depth_order=/' -e 's/bin_count.*resolve=/resolve=/' -e 's/grad_clear.*//' | tr '\n' ' ')"
  done
done
