for cfg in "0 0.35,1" "1 0.35,1" "1 0.5,1" "1 0.7,1" "1 0.85,1" "1 1,1" "1 0.3,0.8,1" "1 0.5,0.9,1"; do
  set -- $cfg
  echo "FACING=$1 BUCKETS=$2: $(DIRT_B200_FACING=$1 DIRT_B200_BUCKETS=$2 python tools/stages.py headline c4 2>&1 | grep -E 'headline|c4' | sed -e 's/clear=.*depth_order=/depth_order=/' -e 's/bin_count.*resolve=/resolve=/' | tr '\n' ' ')"
done
