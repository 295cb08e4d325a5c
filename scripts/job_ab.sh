# A/B on one box: the committed kernels against the working tree's, same bench, alternating.  variants/libacc_head.so is built here
# first (variants/ is git-ignored but travels with gpurun): for f in $(git ls-files acclimate_b200/csrc); do git show HEAD:$f > variants/head_csrc/$(basename $f); done;
# nvcc <the flags of __graft_entry__.py> variants/head_csrc/engine.cu -o variants/libacc_head.so
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --flood-steps 0 --members 0"
for i in 1 2; do
  for L in variants/libacc_head.so acclimate_b200/libacclimate_b200.so $EXTRA_LIBS; do
    ACCLIMATE_B200_LIB=$L timeout 300 $B 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); k=d['roofline']['kernels_ms_per_step']; print('$L', round(d['ms_per_step'],4), {n: round(v,4) for n,v in k.items()})
"
  done
done
