// TensorRT plugin glue for the sm_100a attention kernel (see lightglue_attention_plugin.h).
//
// Behaviour kept from the reference (/root/reference/lightglue_attention_plugin/lightglue_attention_plugin.cpp):
//   name/version/namespace "MHAHeadDim64"/"1"/""                     (:19-20)
//   empty PluginFieldCollection, createPlugin ignores it              (:363-367, :391-404)
//   0-byte serialisation, deserialising ctor ignores its arguments    (:30, :45-47)
//   out dims = Q dims, out type = input-0 type                        (:79-94, :298-310)
//   supportsFormatCombination truth table                             (:272-296)
//   contract violations -> log, cudaDeviceReset(), abort()            (common/checkMacrosPlugin.cpp:118-127)
// Changed: enqueue() is one launcher call (no pad / unpad kernels) and returns non-zero when the
// launcher reports an error (the reference always returned 0, :269); getWorkspaceSize() asks the
// launcher (0 bytes for kHALF) instead of a fixed 5 MiB (:96-112).

#include "lightglue_attention_plugin.h"

#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <exception>

#include "../../../include/lg_attn.h"

using nvinfer1::plugin::LightGlueAttentionPlugin;
using nvinfer1::plugin::LightGlueAttentionPluginCreator;

namespace {

constexpr char const* kPluginName{"MHAHeadDim64"};
constexpr char const* kPluginVersion{"1"};

[[noreturn]] void contract_violation(char const* expr, char const* file, int line) {
    std::fprintf(stderr, "[lightglue_attention_plugin] assertion failed: %s (%s:%d); aborting\n", expr, file, line);
    std::fflush(stderr);
    cudaDeviceReset();
    std::abort();
}

#define LG_PLUGIN_ASSERT(cond)                                      \
    do {                                                            \
        if (!(cond)) contract_violation(#cond, __FILE__, __LINE__); \
    } while (0)

void log_exception(std::exception const& e) noexcept {
    std::fprintf(stderr, "[lightglue_attention_plugin] caught: %s\n", e.what());
}

// The I/O contract, checked identically at configure time and at enqueue time:
// rank-4 [1, 4, n<=2048, 64], K/V agree, O mirrors Q, one common type in {float, half}, linear.
void assert_io_contract(nvinfer1::PluginTensorDesc const& q, nvinfer1::PluginTensorDesc const& k,
                        nvinfer1::PluginTensorDesc const& v, nvinfer1::PluginTensorDesc const& o) {
    nvinfer1::PluginTensorDesc const* all[4] = {&q, &k, &v, &o};
    for (auto const* t : all) {
        LG_PLUGIN_ASSERT(t->dims.nbDims == 4);
        LG_PLUGIN_ASSERT(t->dims.d[0] == ATTENTION_BATCH);
        LG_PLUGIN_ASSERT(t->dims.d[1] == ATTENTION_NUM_HEADS);
        LG_PLUGIN_ASSERT(t->dims.d[2] <= ATTENTION_MAX_SEQ_LEN);
        LG_PLUGIN_ASSERT(t->dims.d[3] == ATTENTION_HEAD_DIM);
        LG_PLUGIN_ASSERT(t->format == nvinfer1::TensorFormat::kLINEAR);
        LG_PLUGIN_ASSERT(t->type == q.type);
    }
    LG_PLUGIN_ASSERT(q.type == nvinfer1::DataType::kFLOAT || q.type == nvinfer1::DataType::kHALF);
    LG_PLUGIN_ASSERT(k.dims.d[2] == v.dims.d[2]);
    LG_PLUGIN_ASSERT(q.dims.d[2] == o.dims.d[2]);
}

int io_mode(nvinfer1::DataType t) { return t == nvinfer1::DataType::kHALF ? LG_ATTN_F16_F16 : LG_ATTN_F32_F32; }

}  // namespace

nvinfer1::PluginFieldCollection LightGlueAttentionPluginCreator::m_plugin_field_collection{};
std::vector<nvinfer1::PluginField> LightGlueAttentionPluginCreator::m_plugin_field_vector;

// ------------------------------------------------------------------------------------ plugin
LightGlueAttentionPlugin::LightGlueAttentionPlugin() = default;
LightGlueAttentionPlugin::LightGlueAttentionPlugin(const void*, size_t) {}
LightGlueAttentionPlugin::~LightGlueAttentionPlugin() { terminate(); }

char const* LightGlueAttentionPlugin::getPluginType() const noexcept { return kPluginName; }
char const* LightGlueAttentionPlugin::getPluginVersion() const noexcept { return kPluginVersion; }
int32_t LightGlueAttentionPlugin::getNbOutputs() const noexcept { return 1; }

int32_t LightGlueAttentionPlugin::initialize() noexcept {
    // Per-device setup (shared-memory opt-in, driver entry point) on whatever device this context
    // runs on; the reference did this once at dlopen for the then-current device only
    // (attention_headdim_64_fp16in_fp16out.cu:242-250).  Not fatal when no device is present
    // (e.g. engine inspection on a CPU host): enqueue re-checks.
    int device = -1;
    if (cudaGetDevice(&device) == cudaSuccess && device >= 0) (void)lg_attn_hd64_init_device(device);
    (void)cudaGetLastError();
    return 0;
}

void LightGlueAttentionPlugin::terminate() noexcept {}
size_t LightGlueAttentionPlugin::getSerializationSize() const noexcept { return 0; }
void LightGlueAttentionPlugin::serialize(void*) const noexcept {}
void LightGlueAttentionPlugin::destroy() noexcept { delete this; }
void LightGlueAttentionPlugin::setPluginNamespace(char const* ns) noexcept { m_plugin_namespace = ns ? ns : ""; }
char const* LightGlueAttentionPlugin::getPluginNamespace() const noexcept { return m_plugin_namespace.c_str(); }

nvinfer1::DataType LightGlueAttentionPlugin::getOutputDataType(int32_t output_index,
                                                               const nvinfer1::DataType* inputs_type,
                                                               int32_t nb_inputs) const noexcept {
    LG_PLUGIN_ASSERT(inputs_type != nullptr && nb_inputs == 3 && output_index == 0);
    return inputs_type[0];
}

void LightGlueAttentionPlugin::attachToContext(cudnnContext*, cublasContext*, nvinfer1::IGpuAllocator*) noexcept {}
void LightGlueAttentionPlugin::detachFromContext() noexcept {}

nvinfer1::IPluginV2DynamicExt* LightGlueAttentionPlugin::clone() const noexcept {
    try {
        auto* copy = new LightGlueAttentionPlugin();
        copy->setPluginNamespace(m_plugin_namespace.c_str());
        copy->initialize();
        return copy;
    } catch (std::exception const& e) {
        log_exception(e);
    }
    return nullptr;
}

nvinfer1::DimsExprs LightGlueAttentionPlugin::getOutputDimensions(int32_t output_index,
                                                                  const nvinfer1::DimsExprs* inputs_dim,
                                                                  int32_t nb_inputs,
                                                                  nvinfer1::IExprBuilder&) noexcept {
    LG_PLUGIN_ASSERT(inputs_dim != nullptr && nb_inputs == 3 && output_index == 0);
    LG_PLUGIN_ASSERT(inputs_dim[0].nbDims == 4 && inputs_dim[1].nbDims == 4 && inputs_dim[2].nbDims == 4);
    return inputs_dim[0];  // O has Q's shape
}

bool LightGlueAttentionPlugin::supportsFormatCombination(int32_t index,
                                                         const nvinfer1::PluginTensorDesc* io,
                                                         int32_t nb_inputs, int32_t nb_outputs) noexcept {
    LG_PLUGIN_ASSERT(io != nullptr && nb_inputs == 3 && nb_outputs == 1 && index >= 0 && index <= 3);
    bool const linear = io[index].format == nvinfer1::TensorFormat::kLINEAR;
    if (index == 0) {
        return linear && (io[0].type == nvinfer1::DataType::kFLOAT || io[0].type == nvinfer1::DataType::kHALF);
    }
    return linear && io[index].type == io[0].type;
}

void LightGlueAttentionPlugin::configurePlugin(const nvinfer1::DynamicPluginTensorDesc* in, int32_t nb_inputs,
                                               const nvinfer1::DynamicPluginTensorDesc* out,
                                               int32_t nb_outputs) noexcept {
    LG_PLUGIN_ASSERT(in != nullptr && out != nullptr && nb_inputs == 3 && nb_outputs == 1);
    // dynamic dims arrive as -1 at build time; -1 <= 2048 holds, like in the reference's checks
    assert_io_contract(in[0].desc, in[1].desc, in[2].desc, out[0].desc);
}

size_t LightGlueAttentionPlugin::getWorkspaceSize(const nvinfer1::PluginTensorDesc* in, int32_t nb_inputs,
                                                  const nvinfer1::PluginTensorDesc* out,
                                                  int32_t nb_outputs) const noexcept {
    LG_PLUGIN_ASSERT(in != nullptr && out != nullptr && nb_inputs == 3 && nb_outputs == 1);
    // TensorRT passes the profile's maximum shapes here.
    int const n_q = in[0].dims.d[2] > 0 ? in[0].dims.d[2] : ATTENTION_MAX_SEQ_LEN;
    int const n_kv = in[1].dims.d[2] > 0 ? in[1].dims.d[2] : ATTENTION_MAX_SEQ_LEN;
    return lg_attn_hd64_workspace_bytes(ATTENTION_BATCH, ATTENTION_NUM_HEADS, n_q, n_kv, io_mode(in[0].type));
}

int32_t LightGlueAttentionPlugin::enqueue(const nvinfer1::PluginTensorDesc* in, const nvinfer1::PluginTensorDesc* out,
                                          const void* const* inputs_ptr, void* const* outputs_ptr, void* workspace,
                                          cudaStream_t stream) noexcept {
    LG_PLUGIN_ASSERT(in != nullptr && out != nullptr && inputs_ptr != nullptr && outputs_ptr != nullptr);
    assert_io_contract(in[0], in[1], in[2], out[0]);
    int const n_q = in[0].dims.d[2];
    int const n_kv = in[1].dims.d[2];
    int const mode = io_mode(in[0].type);
    size_t const ws_bytes = lg_attn_hd64_workspace_bytes(ATTENTION_BATCH, ATTENTION_NUM_HEADS, n_q, n_kv, mode);
    LG_PLUGIN_ASSERT(ws_bytes == 0 || workspace != nullptr);
    int const rc = lg_attn_hd64_launch(inputs_ptr[0], inputs_ptr[1], inputs_ptr[2], outputs_ptr[0], ATTENTION_BATCH,
                                       ATTENTION_NUM_HEADS, n_q, n_kv, mode, workspace, ws_bytes, stream);
    if (rc != 0) {
        std::fprintf(stderr, "[lightglue_attention_plugin] launch failed: %s\n", lg_attn_error_string(rc));
        return rc;
    }
    return 0;
}

// ------------------------------------------------------------------------------------ creator
LightGlueAttentionPluginCreator::LightGlueAttentionPluginCreator() {
    m_plugin_field_vector.clear();  // the ONNX node carries no attributes (lightglue.py:21-26)
    m_plugin_field_collection.nbFields = static_cast<int32_t>(m_plugin_field_vector.size());
    m_plugin_field_collection.fields = m_plugin_field_vector.data();
}

LightGlueAttentionPluginCreator::~LightGlueAttentionPluginCreator() = default;

char const* LightGlueAttentionPluginCreator::getPluginName() const noexcept { return kPluginName; }
char const* LightGlueAttentionPluginCreator::getPluginVersion() const noexcept { return kPluginVersion; }
nvinfer1::PluginFieldCollection const* LightGlueAttentionPluginCreator::getFieldNames() noexcept {
    return &m_plugin_field_collection;
}
void LightGlueAttentionPluginCreator::setPluginNamespace(char const* ns) noexcept { m_plugin_namespace = ns ? ns : ""; }
char const* LightGlueAttentionPluginCreator::getPluginNamespace() const noexcept { return m_plugin_namespace.c_str(); }

nvinfer1::IPluginV2DynamicExt* LightGlueAttentionPluginCreator::createPlugin(char const*,
                                                                             const nvinfer1::PluginFieldCollection*) noexcept {
    try {
        auto* plugin = new LightGlueAttentionPlugin();
        plugin->setPluginNamespace(m_plugin_namespace.c_str());
        plugin->initialize();
        return plugin;
    } catch (std::exception const& e) {
        log_exception(e);
    }
    return nullptr;
}

nvinfer1::IPluginV2DynamicExt* LightGlueAttentionPluginCreator::deserializePlugin(char const*, void const* data,
                                                                                  size_t length) noexcept {
    try {
        auto* plugin = new LightGlueAttentionPlugin(data, length);
        plugin->setPluginNamespace(m_plugin_namespace.c_str());
        plugin->initialize();
        return plugin;
    } catch (std::exception const& e) {
        log_exception(e);
    }
    return nullptr;
}

namespace nvinfer1 {
namespace plugin {
REGISTER_TENSORRT_PLUGIN(LightGlueAttentionPluginCreator);
}  // namespace plugin
}  // namespace nvinfer1
