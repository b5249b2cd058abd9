# GPU box: A/B of throughput-kernel variants.  Libraries are built in the container into csrc/build/variants/<tag>
# (build.py: build(force=True, defs=[...], out=...)); env knobs (LG_ATTN_PERSIST, LG_ATTN_SPLITS) select kernels at run time.
#   bash tools/ab.sh burst|sustained tag [tag ...]      ("main" = the product library)
V=$PWD/lightglue-with-flashattentionv2-tensorrt_b200/csrc/build/variants
mode=$1; shift
for t in "$@"; do
  echo "== $t"
  if [ "$t" = main ]; then lib=""; else lib="LG_ATTN_LIB=$V/$t/liblightglue_attention_plugin.so"; fi
  if [ "$mode" = sustained ]; then env $lib timeout 120 python tools/clock_probe.py 3 2>&1 | grep -A3 "^ours\|^sdpa"
  else env $lib timeout 120 python tools/quick_tf.py 2>&1 | tail -7; fi
  sleep 2
done
