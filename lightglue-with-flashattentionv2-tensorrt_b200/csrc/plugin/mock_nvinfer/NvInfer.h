// MOCK of the small slice of TensorRT's <NvInfer.h> that an IPluginV2DynamicExt plugin touches.
// Used ONLY when the real TensorRT headers are absent (they are, in this image and on the GPU box:
// SURVEY.md §0.1).  It lets the plugin translation unit compile and lets a mock host
// (mock_trt_host.cpp) drive the same call sequence TensorRT does:
//   registry lookup -> createPlugin -> getOutputDimensions -> supportsFormatCombination ->
//   configurePlugin -> getWorkspaceSize -> clone -> enqueue -> destroy.
// Interface names / signatures follow TensorRT 8.5-10.x's public API so that building against the
// real header needs no source change (the build picks the real one when NvInfer.h is found).
#ifndef LG_MOCK_NVINFER_H_
#define LG_MOCK_NVINFER_H_
#define LG_ATTN_MOCK_TENSORRT 1

#include <cuda_runtime_api.h>
#include <stddef.h>
#include <stdint.h>

struct cudnnContext;
struct cublasContext;

namespace nvinfer1 {

using AsciiChar = char;

enum class DataType : int32_t { kFLOAT = 0, kHALF = 1, kINT8 = 2, kINT32 = 3, kBOOL = 4 };
enum class TensorFormat : int32_t { kLINEAR = 0, kCHW2 = 1, kHWC8 = 2, kCHW4 = 3, kCHW16 = 4, kCHW32 = 5 };
using PluginFormat = TensorFormat;

struct Dims {
    static constexpr int32_t MAX_DIMS = 8;
    int32_t nbDims;
    int32_t d[MAX_DIMS];
};

class IDimensionExpr {
public:
    virtual bool isConstant() const noexcept = 0;
    virtual int32_t getConstantValue() const noexcept = 0;
protected:
    virtual ~IDimensionExpr() noexcept = default;
};

struct DimsExprs {
    int32_t nbDims;
    IDimensionExpr const* d[Dims::MAX_DIMS];
};

class IExprBuilder {
public:
    virtual IDimensionExpr const* constant(int32_t value) noexcept = 0;
protected:
    virtual ~IExprBuilder() noexcept = default;
};

struct PluginTensorDesc {
    Dims dims;
    DataType type;
    TensorFormat format;
    float scale;
};

struct DynamicPluginTensorDesc {
    PluginTensorDesc desc;
    Dims min;
    Dims max;
};

enum class PluginFieldType : int32_t { kFLOAT16 = 0, kFLOAT32 = 1, kFLOAT64 = 2, kINT8 = 3, kINT16 = 4, kINT32 = 5, kCHAR = 6, kDIMS = 7, kUNKNOWN = 8 };

struct PluginField {
    AsciiChar const* name;
    void const* data;
    PluginFieldType type;
    int32_t length;
};

struct PluginFieldCollection {
    int32_t nbFields;
    PluginField const* fields;
};

class IGpuAllocator;

class IPluginV2 {
public:
    virtual AsciiChar const* getPluginType() const noexcept = 0;
    virtual AsciiChar const* getPluginVersion() const noexcept = 0;
    virtual int32_t getNbOutputs() const noexcept = 0;
    virtual int32_t initialize() noexcept = 0;
    virtual void terminate() noexcept = 0;
    virtual size_t getSerializationSize() const noexcept = 0;
    virtual void serialize(void* buffer) const noexcept = 0;
    virtual void destroy() noexcept = 0;
    virtual void setPluginNamespace(AsciiChar const* ns) noexcept = 0;
    virtual AsciiChar const* getPluginNamespace() const noexcept = 0;
    virtual ~IPluginV2() noexcept = default;
};

class IPluginV2Ext : public IPluginV2 {
public:
    virtual DataType getOutputDataType(int32_t index, DataType const* inputTypes, int32_t nbInputs) const noexcept = 0;
    virtual void attachToContext(cudnnContext*, cublasContext*, IGpuAllocator*) noexcept {}
    virtual void detachFromContext() noexcept {}
};

class IPluginV2DynamicExt : public IPluginV2Ext {
public:
    virtual IPluginV2DynamicExt* clone() const noexcept = 0;
    virtual DimsExprs getOutputDimensions(int32_t outputIndex, DimsExprs const* inputs, int32_t nbInputs,
                                          IExprBuilder& exprBuilder) noexcept = 0;
    virtual bool supportsFormatCombination(int32_t pos, PluginTensorDesc const* inOut, int32_t nbInputs,
                                           int32_t nbOutputs) noexcept = 0;
    virtual void configurePlugin(DynamicPluginTensorDesc const* in, int32_t nbInputs,
                                 DynamicPluginTensorDesc const* out, int32_t nbOutputs) noexcept = 0;
    virtual size_t getWorkspaceSize(PluginTensorDesc const* inputs, int32_t nbInputs,
                                    PluginTensorDesc const* outputs, int32_t nbOutputs) const noexcept = 0;
    virtual int32_t enqueue(PluginTensorDesc const* inputDesc, PluginTensorDesc const* outputDesc,
                            void const* const* inputs, void* const* outputs, void* workspace,
                            cudaStream_t stream) noexcept = 0;
};

class IPluginCreator {
public:
    virtual AsciiChar const* getPluginName() const noexcept = 0;
    virtual AsciiChar const* getPluginVersion() const noexcept = 0;
    virtual PluginFieldCollection const* getFieldNames() noexcept = 0;
    virtual IPluginV2* createPlugin(AsciiChar const* name, PluginFieldCollection const* fc) noexcept = 0;
    virtual IPluginV2* deserializePlugin(AsciiChar const* name, void const* data, size_t length) noexcept = 0;
    virtual void setPluginNamespace(AsciiChar const* ns) noexcept = 0;
    virtual AsciiChar const* getPluginNamespace() const noexcept = 0;
    virtual ~IPluginCreator() = default;
};

class IPluginRegistry {
public:
    virtual bool registerCreator(IPluginCreator& creator, AsciiChar const* pluginNamespace) noexcept = 0;
    virtual IPluginCreator* getPluginCreator(AsciiChar const* type, AsciiChar const* version,
                                             AsciiChar const* pluginNamespace = "") noexcept = 0;
    virtual IPluginCreator* const* getPluginCreatorList(int32_t* numCreators) const noexcept = 0;
protected:
    virtual ~IPluginRegistry() noexcept = default;
};

template <typename T>
class PluginRegistrar {
public:
    PluginRegistrar();
private:
    T instance{};
};

}  // namespace nvinfer1

extern "C" nvinfer1::IPluginRegistry* getPluginRegistry() noexcept;

namespace nvinfer1 {
template <typename T>
PluginRegistrar<T>::PluginRegistrar() {
    getPluginRegistry()->registerCreator(instance, "");
}
}  // namespace nvinfer1

#define REGISTER_TENSORRT_PLUGIN(name) \
    static nvinfer1::PluginRegistrar<name> pluginRegistrar##name {}

#endif  // LG_MOCK_NVINFER_H_
