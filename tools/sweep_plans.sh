# development aid: times bucket plans ("a,b;c,1": ';' ends a launch, ',' a bucket inside it) on the headline and c4 workloads
for plan in "0.35;1" "0.1,0.35;1" "0.05,0.15,0.35;1" "0.1,0.2,0.35;1" "0.35;0.6,1" "0.1,0.35;0.6,1" "0.1,0.2,0.35;0.5,0.7,1" "0.1,0.25,0.45;1" "0.05,0.1,0.2,0.35;0.5,0.7,1" "0.1,0.2,0.3,0.4,0.5;1"; do
  echo "BUCKETS=$plan: $(DIRT_B200_BUCKETS=$plan python tools/stages.py headline c4 2>&1 | grep -E 'headline|c4' | sed -e 's/clear=.*depth_order=/depth_order=/' -e 's/bin_count.*resolve=/resolve=/' -e 's/grad_clear.*//' | tr '\n' ' ')"
done
