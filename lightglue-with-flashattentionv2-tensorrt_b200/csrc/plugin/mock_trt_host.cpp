// Mock TensorRT host (built only with the mock NvInfer.h): a plugin registry plus C entry points
// that walk a plugin through the call sequence TensorRT uses, so the plugin boundary is tested even
// where TensorRT itself is not installed.  Sequence mirrors SURVEY.md §3.1/§3.2:
//   getPluginRegistry()->getPluginCreator("MHAHeadDim64","1","") -> createPlugin (or deserializePlugin)
//   -> getNbOutputs / getOutputDataType / getOutputDimensions -> supportsFormatCombination (all 4 slots)
//   -> configurePlugin (dynamic dims, min/max) -> getWorkspaceSize (max shapes) -> clone (per context)
//   -> enqueue(descs, pointers, workspace, stream) -> destroy.
// Test infrastructure only; never part of a product call path.
#include <NvInfer.h>

#ifdef LG_ATTN_MOCK_TENSORRT

#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace {

class MockRegistry final : public nvinfer1::IPluginRegistry {
public:
    bool registerCreator(nvinfer1::IPluginCreator& creator, char const* ns) noexcept override {
        creator.setPluginNamespace(ns ? ns : "");
        creators_.push_back(&creator);
        return true;
    }
    nvinfer1::IPluginCreator* getPluginCreator(char const* type, char const* version, char const* ns) noexcept override {
        for (auto* c : creators_) {
            if (std::strcmp(c->getPluginName(), type) == 0 && std::strcmp(c->getPluginVersion(), version) == 0 &&
                std::strcmp(c->getPluginNamespace(), ns ? ns : "") == 0)
                return c;
        }
        return nullptr;
    }
    nvinfer1::IPluginCreator* const* getPluginCreatorList(int32_t* n) const noexcept override {
        *n = static_cast<int32_t>(creators_.size());
        return creators_.data();
    }

private:
    std::vector<nvinfer1::IPluginCreator*> creators_;
};

class ConstExpr final : public nvinfer1::IDimensionExpr {
public:
    explicit ConstExpr(int32_t v) : v_(v) {}
    bool isConstant() const noexcept override { return v_ >= 0; }
    int32_t getConstantValue() const noexcept override { return v_; }

private:
    int32_t v_;
};

class MockExprBuilder final : public nvinfer1::IExprBuilder {
public:
    nvinfer1::IDimensionExpr const* constant(int32_t v) noexcept override {
        pool_.emplace_back(new ConstExpr(v));
        return pool_.back();
    }
    ~MockExprBuilder() override {
        for (auto* p : pool_) delete p;
    }

private:
    std::vector<ConstExpr*> pool_;
};

nvinfer1::PluginTensorDesc make_desc(int rows, nvinfer1::DataType t) {
    nvinfer1::PluginTensorDesc d{};
    d.dims.nbDims = 4;
    d.dims.d[0] = 1;
    d.dims.d[1] = 4;
    d.dims.d[2] = rows;
    d.dims.d[3] = 64;
    d.type = t;
    d.format = nvinfer1::TensorFormat::kLINEAR;
    d.scale = 1.f;
    return d;
}

nvinfer1::IPluginV2DynamicExt* make_plugin(bool via_deserialize) {
    auto* creator = getPluginRegistry()->getPluginCreator("MHAHeadDim64", "1", "");
    if (creator == nullptr) return nullptr;
    nvinfer1::PluginFieldCollection const* fc = creator->getFieldNames();
    nvinfer1::IPluginV2* p = via_deserialize ? creator->deserializePlugin("attn", nullptr, 0)
                                             : creator->createPlugin("attn", fc);
    return static_cast<nvinfer1::IPluginV2DynamicExt*>(p);
}

}  // namespace

extern "C" nvinfer1::IPluginRegistry* getPluginRegistry() noexcept {
    static MockRegistry registry;
    return &registry;
}

extern "C" {

// Describe the registered plugin without touching a GPU. Returns bytes written (JSON), <0 on error.
int lg_mock_trt_describe(char* buf, size_t len) {
    auto* creator = getPluginRegistry()->getPluginCreator("MHAHeadDim64", "1", "");
    if (creator == nullptr) return -1;
    int32_t n_creators = 0;
    getPluginRegistry()->getPluginCreatorList(&n_creators);
    auto* plugin = make_plugin(false);
    if (plugin == nullptr) return -2;

    // supportsFormatCombination truth table over {kFLOAT,kHALF,kINT8} x {kLINEAR,kCHW4} for slot 0, and
    // for slots 1-3 given slot 0 = half/linear
    std::string table;
    nvinfer1::DataType const types[3] = {nvinfer1::DataType::kFLOAT, nvinfer1::DataType::kHALF, nvinfer1::DataType::kINT8};
    nvinfer1::TensorFormat const fmts[2] = {nvinfer1::TensorFormat::kLINEAR, nvinfer1::TensorFormat::kCHW4};
    for (int slot = 0; slot < 4; ++slot) {
        for (auto t : types) {
            for (auto f : fmts) {
                nvinfer1::PluginTensorDesc io[4] = {make_desc(8, nvinfer1::DataType::kHALF), make_desc(8, nvinfer1::DataType::kHALF),
                                                    make_desc(8, nvinfer1::DataType::kHALF), make_desc(8, nvinfer1::DataType::kHALF)};
                io[slot].type = t;
                io[slot].format = f;
                table += plugin->supportsFormatCombination(slot, io, 3, 1) ? '1' : '0';
            }
        }
    }
    // slots 1-3 must follow slot 0 when slot 0 is float
    std::string follow;
    for (int slot = 1; slot < 4; ++slot) {
        nvinfer1::PluginTensorDesc io[4] = {make_desc(8, nvinfer1::DataType::kFLOAT), make_desc(8, nvinfer1::DataType::kFLOAT),
                                            make_desc(8, nvinfer1::DataType::kFLOAT), make_desc(8, nvinfer1::DataType::kFLOAT)};
        follow += plugin->supportsFormatCombination(slot, io, 3, 1) ? '1' : '0';
        io[slot].type = nvinfer1::DataType::kHALF;
        follow += plugin->supportsFormatCombination(slot, io, 3, 1) ? '1' : '0';
    }

    MockExprBuilder eb;
    nvinfer1::DimsExprs in_dims[3];
    int const rows[3] = {777, 1000, 1000};
    for (int i = 0; i < 3; ++i) {
        in_dims[i].nbDims = 4;
        in_dims[i].d[0] = eb.constant(1);
        in_dims[i].d[1] = eb.constant(4);
        in_dims[i].d[2] = eb.constant(rows[i]);
        in_dims[i].d[3] = eb.constant(64);
    }
    nvinfer1::DimsExprs od = plugin->getOutputDimensions(0, in_dims, 3, eb);
    nvinfer1::DataType in_t[3] = {nvinfer1::DataType::kHALF, nvinfer1::DataType::kHALF, nvinfer1::DataType::kHALF};
    int const out_t_half = static_cast<int>(plugin->getOutputDataType(0, in_t, 3));
    in_t[0] = in_t[1] = in_t[2] = nvinfer1::DataType::kFLOAT;
    int const out_t_float = static_cast<int>(plugin->getOutputDataType(0, in_t, 3));

    nvinfer1::PluginTensorDesc hin[3] = {make_desc(2048, nvinfer1::DataType::kHALF), make_desc(2048, nvinfer1::DataType::kHALF),
                                         make_desc(2048, nvinfer1::DataType::kHALF)};
    nvinfer1::PluginTensorDesc hout = make_desc(2048, nvinfer1::DataType::kHALF);
    size_t const ws_half = plugin->getWorkspaceSize(hin, 3, &hout, 1);
    nvinfer1::PluginTensorDesc fin[3] = {make_desc(2048, nvinfer1::DataType::kFLOAT), make_desc(2048, nvinfer1::DataType::kFLOAT),
                                         make_desc(2048, nvinfer1::DataType::kFLOAT)};
    nvinfer1::PluginTensorDesc fout = make_desc(2048, nvinfer1::DataType::kFLOAT);
    size_t const ws_float = plugin->getWorkspaceSize(fin, 3, &fout, 1);

    plugin->setPluginNamespace("ns_probe");
    auto* cl = plugin->clone();
    std::string clone_ns = cl ? cl->getPluginNamespace() : "<null>";
    char ser[4] = {0x5a, 0x5a, 0x5a, 0x5a};
    plugin->serialize(ser);
    bool const ser_untouched = ser[0] == 0x5a && ser[3] == 0x5a;

    int n = std::snprintf(
        buf, len,
        "{\"name\": \"%s\", \"version\": \"%s\", \"namespace\": \"%s\", \"n_creators\": %d, \"nb_fields\": %d, "
        "\"fields_null\": %s, \"type\": \"%s\", \"plugin_version\": \"%s\", \"nb_outputs\": %d, \"serialization_size\": %zu, "
        "\"serialize_untouched\": %s, \"format_table\": \"%s\", \"format_follow_float\": \"%s\", \"out_nb_dims\": %d, "
        "\"out_dims\": [%d, %d, %d, %d], \"out_type_half\": %d, \"out_type_float\": %d, \"ws_half\": %zu, \"ws_float\": %zu, "
        "\"clone_namespace\": \"%s\"}",
        creator->getPluginName(), creator->getPluginVersion(), creator->getPluginNamespace(), n_creators,
        creator->getFieldNames()->nbFields, creator->getFieldNames()->nbFields == 0 ? "true" : "false", plugin->getPluginType(),
        plugin->getPluginVersion(), plugin->getNbOutputs(), plugin->getSerializationSize(), ser_untouched ? "true" : "false",
        table.c_str(), follow.c_str(), od.nbDims, od.d[0]->getConstantValue(), od.d[1]->getConstantValue(),
        od.d[2]->getConstantValue(), od.d[3]->getConstantValue(), out_t_half, out_t_float, ws_half, ws_float, clone_ns.c_str());
    if (cl) cl->destroy();
    plugin->destroy();
    return n;
}

// Full TensorRT-style walk ending in enqueue() on `stream`. Device pointers. Returns enqueue's code,
// or a negative number when the host-side walk itself fails.
int lg_mock_trt_enqueue(const void* q, const void* k, const void* v, void* o, int n_q, int n_kv, int is_float,
                        void* workspace, size_t workspace_bytes, void* stream, int via_deserialize, int use_clone) {
    auto* built = make_plugin(via_deserialize != 0);
    if (built == nullptr) return -1;
    nvinfer1::DataType const t = is_float ? nvinfer1::DataType::kFLOAT : nvinfer1::DataType::kHALF;
    if (built->getNbOutputs() != 1) return -2;

    // build phase: format negotiation on every slot
    nvinfer1::PluginTensorDesc io[4] = {make_desc(n_q, t), make_desc(n_kv, t), make_desc(n_kv, t), make_desc(n_q, t)};
    for (int slot = 0; slot < 4; ++slot)
        if (!built->supportsFormatCombination(slot, io, 3, 1)) return -3;
    nvinfer1::DataType in_t[3] = {t, t, t};
    if (built->getOutputDataType(0, in_t, 3) != t) return -4;

    // dynamic-shape configure: desc dims -1 on the keypoint axis, profile min 1 / max 2048 (steps.txt:152)
    nvinfer1::DynamicPluginTensorDesc din[3], dout[1];
    for (int i = 0; i < 3; ++i) {
        din[i].desc = make_desc(-1, t);
        din[i].min = make_desc(1, t).dims;
        din[i].max = make_desc(2048, t).dims;
    }
    dout[0].desc = make_desc(-1, t);
    dout[0].min = make_desc(1, t).dims;
    dout[0].max = make_desc(2048, t).dims;
    built->configurePlugin(din, 3, dout, 1);

    nvinfer1::PluginTensorDesc maxin[3] = {make_desc(2048, t), make_desc(2048, t), make_desc(2048, t)};
    nvinfer1::PluginTensorDesc maxout = make_desc(2048, t);
    size_t const ws_need = built->getWorkspaceSize(maxin, 3, &maxout, 1);
    if (ws_need > workspace_bytes) {
        built->destroy();
        return -5;
    }

    // runtime phase: TensorRT clones the plugin per execution context
    nvinfer1::IPluginV2DynamicExt* run = use_clone ? built->clone() : built;
    if (run == nullptr) return -6;
    void const* ins[3] = {q, k, v};
    void* outs[1] = {o};
    int const rc = run->enqueue(io, io + 3, ins, outs, workspace, static_cast<cudaStream_t>(stream));
    if (use_clone) run->destroy();
    built->destroy();
    return rc;
}

// workspace size TensorRT would allocate for the 1x4x2048x64 max profile
size_t lg_mock_trt_workspace_bytes(int is_float) {
    auto* p = make_plugin(false);
    if (p == nullptr) return 0;
    nvinfer1::DataType const t = is_float ? nvinfer1::DataType::kFLOAT : nvinfer1::DataType::kHALF;
    nvinfer1::PluginTensorDesc maxin[3] = {make_desc(2048, t), make_desc(2048, t), make_desc(2048, t)};
    nvinfer1::PluginTensorDesc maxout = make_desc(2048, t);
    size_t const ws = p->getWorkspaceSize(maxin, 3, &maxout, 1);
    p->destroy();
    return ws;
}

}  // extern "C"

#endif  // LG_ATTN_MOCK_TENSORRT
