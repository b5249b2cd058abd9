// TensorRT plugin "MHAHeadDim64" / version "1" / namespace "" — B200 (sm_100a) build.
//
// Same public shape as the reference plugin
// (/root/reference/lightglue_attention_plugin/lightglue_attention_plugin.h:28-159): an
// IPluginV2DynamicExt with 3 inputs (Q, K, V) and 1 output, an IPluginCreator with an EMPTY
// PluginFieldCollection, 0-byte serialisation, registration by static initialiser.  What changed is
// behind enqueue(): one call into the extern "C" launcher of include/lg_attn.h instead of the
// pad -> CUTLASS kernel -> unpad sequence.
#ifndef LG_LIGHTGLUE_ATTENTION_PLUGIN_H_
#define LG_LIGHTGLUE_ATTENTION_PLUGIN_H_

#include <cuda_runtime_api.h>

#include <string>
#include <vector>

#include <NvInfer.h>
#include <NvInferPlugin.h>

// limits the reference asserts on (lightglue_attention_plugin.h:19-22); kept so that engines built
// from the same ONNX / shape profiles are validated identically
#define ATTENTION_BATCH 1
#define ATTENTION_NUM_HEADS 4
#define ATTENTION_MAX_SEQ_LEN 2048
#define ATTENTION_HEAD_DIM 64

namespace nvinfer1 {
namespace plugin {

class LightGlueAttentionPlugin : public nvinfer1::IPluginV2DynamicExt {
public:
    LightGlueAttentionPlugin();
    LightGlueAttentionPlugin(const void* serialized_data, size_t serialized_length);
    ~LightGlueAttentionPlugin() override;

    // IPluginV2
    char const* getPluginType() const noexcept override;
    char const* getPluginVersion() const noexcept override;
    int32_t getNbOutputs() const noexcept override;
    int32_t initialize() noexcept override;
    void terminate() noexcept override;
    size_t getSerializationSize() const noexcept override;
    void serialize(void* buffer) const noexcept override;
    void destroy() noexcept override;
    void setPluginNamespace(char const* plugin_namespace) noexcept override;
    char const* getPluginNamespace() const noexcept override;

    // IPluginV2Ext
    nvinfer1::DataType getOutputDataType(int32_t output_index, const nvinfer1::DataType* inputs_type,
                                         int32_t nb_inputs) const noexcept override;
    void attachToContext(cudnnContext* cudnn_context, cublasContext* cublas_context,
                         nvinfer1::IGpuAllocator* gpu_allocator) noexcept override;
    void detachFromContext() noexcept override;

    // IPluginV2DynamicExt
    nvinfer1::IPluginV2DynamicExt* clone() const noexcept override;
    nvinfer1::DimsExprs getOutputDimensions(int32_t output_index, const nvinfer1::DimsExprs* inputs_dim,
                                            int32_t nb_inputs,
                                            nvinfer1::IExprBuilder& expr_builder) noexcept override;
    bool supportsFormatCombination(int32_t index, const nvinfer1::PluginTensorDesc* inputs_and_outputs_desc,
                                   int32_t nb_inputs, int32_t nb_outputs) noexcept override;
    void configurePlugin(const nvinfer1::DynamicPluginTensorDesc* inputs_desc, int32_t nb_inputs,
                         const nvinfer1::DynamicPluginTensorDesc* outputs_desc,
                         int32_t nb_outputs) noexcept override;
    size_t getWorkspaceSize(const nvinfer1::PluginTensorDesc* inputs_desc, int32_t nb_inputs,
                            const nvinfer1::PluginTensorDesc* outputs_desc,
                            int32_t nb_outputs) const noexcept override;
    int32_t enqueue(const nvinfer1::PluginTensorDesc* inputs_desc, const nvinfer1::PluginTensorDesc* outputs_desc,
                    const void* const* inputs_ptr, void* const* outputs_ptr, void* workspace,
                    cudaStream_t stream) noexcept override;

private:
    std::string m_plugin_namespace;
};

class LightGlueAttentionPluginCreator : public nvinfer1::IPluginCreator {
public:
    LightGlueAttentionPluginCreator();
    ~LightGlueAttentionPluginCreator() override;

    char const* getPluginName() const noexcept override;
    char const* getPluginVersion() const noexcept override;
    nvinfer1::PluginFieldCollection const* getFieldNames() noexcept override;
    nvinfer1::IPluginV2DynamicExt* createPlugin(char const* name,
                                                const nvinfer1::PluginFieldCollection* fc) noexcept override;
    nvinfer1::IPluginV2DynamicExt* deserializePlugin(char const* name, void const* serialized_data,
                                                     size_t serialized_length) noexcept override;
    void setPluginNamespace(char const* plugin_namespace) noexcept override;
    char const* getPluginNamespace() const noexcept override;

private:
    static nvinfer1::PluginFieldCollection m_plugin_field_collection;
    static std::vector<nvinfer1::PluginField> m_plugin_field_vector;
    std::string m_plugin_namespace;
};

}  // namespace plugin
}  // namespace nvinfer1

#endif  // LG_LIGHTGLUE_ATTENTION_PLUGIN_H_
