# GPU box: A/B of throughput-kernel variants: libraries under csrc/build/variants/<tag>, env knobs per run
V=$PWD/lightglue-with-flashattentionv2-tensorrt_b200/csrc/build/variants
run() { echo "== $*"; env "$@" timeout 120 python tools/quick_tf.py 2>&1 | tail -7; sleep 2; }
{
run LG_ATTN_PERSIST=1
run LG_ATTN_PERSIST=1 LG_ATTN_LIB=$V/probe/liblightglue_attention_plugin.so
run LG_ATTN_PERSIST=0
run LG_ATTN_PERSIST=1 LG_ATTN_LIB=$V/probe/liblightglue_attention_plugin.so
} > gpurun_out/ab.txt 2>&1
cat gpurun_out/ab.txt
