// TEST INFRASTRUCTURE (oracle/). Stand-in for the part of DALI's operator API that
// /root/reference/packages/dali_pipeline_framework/ext_impl/DrawGaussians.{h,cc} use: OpSpec arguments, Workspace
// inputs/outputs as lists of per-sample tensors, the thread pool, schema/registration macros (no-ops).
// Written for this repo (not DALI code); the arithmetic under test is entirely in the reference's .cc file.
#pragma once
#include <cstdint>
#include <functional>
#include <map>
#include <string>
#include <type_traits>
#include <vector>

#include "dali/core/error_handling.h"
#include "dali/pipeline/data/types.h"

namespace dali {

template <int N = -1>
class TensorShape {
   public:
    TensorShape() = default;
    template <typename... Ts, typename = std::enable_if_t<(std::is_integral<Ts>::value && ...)>>
    TensorShape(Ts... v) : d_{static_cast<int64_t>(v)...} {}
    size_t size() const { return d_.size(); }
    int64_t operator[](size_t i) const { return d_[i]; }

   private:
    std::vector<int64_t> d_;
};

template <int N = -1>
class TensorListShape {
   public:
    TensorListShape() = default;
    explicit TensorListShape(const std::vector<TensorShape<N>>& s) : s_(s) {}
    int num_samples() const { return (int)s_.size(); }
    const TensorShape<N>& operator[](size_t i) const { return s_[i]; }
    void push_back(const TensorShape<N>& t) { s_.push_back(t); }

   private:
    std::vector<TensorShape<N>> s_;
};

struct OutputDesc {
    TensorListShape<> shape;
    DALIDataType type = DALI_NO_TYPE;
};

// A batch of per-sample tensors: shapes + raw pointers (owned by the driver).
class TensorList {
   public:
    TensorListShape<> shapes;
    std::vector<void*> data;
    DALIDataType dtype = DALI_NO_TYPE;
    const TensorListShape<>& shape() const { return shapes; }  // by reference: callers keep references into it
    DALIDataType type() const { return dtype; }
    const void* raw_tensor(int s) const { return data[s]; }
    void* raw_mutable_tensor(int s) { return data[s]; }
};

class ThreadPool {
   public:
    int NumThreads() const { return 1; }
    void AddWork(std::function<void(int)> f) { work_.push_back(std::move(f)); }
    void RunAll() {
        for (auto& f : work_) f(0);
        work_.clear();
    }

   private:
    std::vector<std::function<void(int)>> work_;
};

class Workspace {
   public:
    std::vector<TensorList> inputs, outputs;
    ThreadPool pool;
    template <typename Backend>
    const TensorList& Input(int i) const { return inputs[i]; }
    template <typename Backend>
    TensorList& Output(int i) { return outputs[i]; }
    ThreadPool& GetThreadPool() { return pool; }
};

class OpSpec {
   public:
    std::map<std::string, std::vector<float>> float_vecs;
    std::map<std::string, float> floats;
    template <typename T>
    std::vector<T> GetRepeatedArgument(const std::string& name) const {
        const auto& v = float_vecs.at(name);
        return std::vector<T>(v.begin(), v.end());
    }
    template <typename T>
    T GetArgument(const std::string& name) const { return static_cast<T>(floats.at(name)); }
};

template <typename Backend>
class Operator {
   public:
    explicit Operator(const OpSpec&) {}
    virtual ~Operator() {}
    bool Setup(std::vector<OutputDesc>& d, const Workspace& ws) { return SetupImpl(d, ws); }
    void Run(Workspace& ws) { RunImpl(ws); }

   protected:
    virtual bool SetupImpl(std::vector<OutputDesc>& output_desc, const Workspace& ws) = 0;
    virtual void RunImpl(Workspace& ws) = 0;
};

struct SchemaStub {
    SchemaStub& DocStr(const char*) { return *this; }
    SchemaStub& NumInput(int) { return *this; }
    SchemaStub& NumOutput(int) { return *this; }
    SchemaStub& AddArg(const char*, const char*, DALIDataType) { return *this; }
};

}  // namespace dali

#define DALI_REGISTER_OPERATOR(name, type, backend) static_assert(true, "registration is a no-op in the stub")
#define DALI_SCHEMA(name) static ::dali::SchemaStub dali_schema_stub_##name = ::dali::SchemaStub()
