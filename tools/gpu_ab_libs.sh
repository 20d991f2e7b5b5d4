# A/B of experiment builds of the library on one box: for each conv-tasnet_b200/libctn_b200_<tag>.so named in $@
# (and the product build) run the command in $CMD; the swap happens only in the box's scratch copy
set -x
export CTN_ALLOW_EXPERIMENT_BUILD=1   # the Python layer refuses non-product operand formats otherwise
mkdir -p gpurun_out
CMD=${CMD:-python tools/bench_configs.py cfg2 cfg5}
cp conv-tasnet_b200/libctn_b200.so /tmp/product.so
for round in 1 2; do
  for tag in product "$@"; do
    if [ $tag = product ]; then cp /tmp/product.so conv-tasnet_b200/libctn_b200.so; else cp conv-tasnet_b200/libctn_b200_$tag.so conv-tasnet_b200/libctn_b200.so; fi
    echo "== $tag (round $round)" >> gpurun_out/ab_libs.log
    timeout 200 $CMD 2>/dev/null | cut -c1-150 >> gpurun_out/ab_libs.log
  done
done
cp /tmp/product.so conv-tasnet_b200/libctn_b200.so
cat gpurun_out/ab_libs.log
