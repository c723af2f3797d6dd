// UNCOMPILED: this file has never met a compiler (no TensorFlow on either machine); it earns no coverage claim.
// TensorFlow custom-op shim over the C-ABI of libkgcnn_b200.so (include/kgcnn_b200.h).
// NOT built in this repository's environment: TensorFlow is not installable on the build or the GPU machine
// (SURVEY.md F2).  Build on a machine with TensorFlow >= 2.9 (reference pin: setup.py:22) via tf_shim/build.sh.
// Each OpKernel reads the op's CUDA stream and forwards raw device pointers; outputs are allocated by TF.
#include "tensorflow/core/framework/op.h"
#include "tensorflow/core/framework/op_kernel.h"
#include "tensorflow/core/framework/shape_inference.h"
#include "tensorflow/core/platform/stream_executor.h"

#include "../include/kgcnn_b200.h"

using namespace tensorflow;  // NOLINT

namespace {
kgcnn_stream_t StreamOf(OpKernelContext* ctx) {
  return reinterpret_cast<kgcnn_stream_t>(
      ctx->op_device_context()->stream()->platform_specific_handle().stream);
}
#define KGCNN_OK_OR_RETURN(ctx, call)                                                        \
  do {                                                                                       \
    int rc__ = (call);                                                                       \
    OP_REQUIRES(ctx, rc__ == 0, errors::Internal(#call, ": ", kgcnn_error_string(rc__)));    \
  } while (0)
}  // namespace

// ---- a1: partition_row_indexing (kgcnn/ops/partition.py:97-161) ------------------------------------
REGISTER_OP("KgcnnShiftIndex")
    .Input("index: int64").Input("node_row_splits: int64").Input("edge_row_splits: int64")
    .Output("index_batch_i32: int32")
    .SetShapeFn([](shape_inference::InferenceContext* c) {
      c->set_output(0, c->MakeShape({c->Dim(c->input(0), 1), c->Dim(c->input(0), 0)}));
      return OkStatus();
    });

class KgcnnShiftIndexOp : public OpKernel {
 public:
  explicit KgcnnShiftIndexOp(OpKernelConstruction* c) : OpKernel(c) {}
  void Compute(OpKernelContext* ctx) override {
    const Tensor& idx = ctx->input(0);
    const Tensor& ns = ctx->input(1);
    const Tensor& es = ctx->input(2);
    const int64_t E = idx.dim_size(0);
    const int K = static_cast<int>(idx.dim_size(1));
    Tensor* out = nullptr;
    OP_REQUIRES_OK(ctx, ctx->allocate_output(0, TensorShape({K, E}), &out));
    KGCNN_OK_OR_RETURN(ctx, kgcnn_shift_index(idx.flat<int64_t>().data(), E, K, ns.flat<int64_t>().data(),
                                               es.flat<int64_t>().data(), es.dim_size(0) - 1, nullptr,
                                               out->flat<int32>().data(), StreamOf(ctx)));
  }
};
REGISTER_KERNEL_BUILDER(Name("KgcnnShiftIndex").Device(DEVICE_GPU), KgcnnShiftIndexOp);

// ---- plan: CSR row pointer + stable permutation (kgcnn/layers/pooling.py:63-68) --------------------------
REGISTER_OP("KgcnnPlanBuild")
    .Input("key: int32").Attr("num_targets: int").Attr("assume_sorted: bool = false")
    .Output("rowptr: int32").Output("perm: int32")
    .SetShapeFn([](shape_inference::InferenceContext* c) {
      int64_t n;
      TF_RETURN_IF_ERROR(c->GetAttr("num_targets", &n));
      c->set_output(0, c->Vector(n + 1));
      c->set_output(1, c->input(0));
      return OkStatus();
    });

class KgcnnPlanBuildOp : public OpKernel {
 public:
  explicit KgcnnPlanBuildOp(OpKernelConstruction* c) : OpKernel(c) {
    OP_REQUIRES_OK(c, c->GetAttr("num_targets", &n_));
    OP_REQUIRES_OK(c, c->GetAttr("assume_sorted", &sorted_));
  }
  void Compute(OpKernelContext* ctx) override {
    const Tensor& key = ctx->input(0);
    const int64_t E = key.dim_size(0);
    Tensor *rowptr = nullptr, *perm = nullptr, ws;
    OP_REQUIRES_OK(ctx, ctx->allocate_output(0, TensorShape({n_ + 1}), &rowptr));
    OP_REQUIRES_OK(ctx, ctx->allocate_output(1, TensorShape({E}), &perm));
    const size_t bytes = kgcnn_plan_workspace_bytes(E, n_);
    OP_REQUIRES_OK(ctx, ctx->allocate_temp(DT_INT8, TensorShape({static_cast<int64_t>(bytes)}), &ws));
    KGCNN_OK_OR_RETURN(ctx, kgcnn_plan_build(key.flat<int32>().data(), 1, E, n_, sorted_ ? 1 : 0,
                                              rowptr->flat<int32>().data(), perm->flat<int32>().data(),
                                              ws.flat<int8>().data(), bytes, StreamOf(ctx)));
  }
 private:
  int64_t n_;
  bool sorted_;
};
REGISTER_KERNEL_BUILDER(Name("KgcnnPlanBuild").Device(DEVICE_GPU), KgcnnPlanBuildOp);

// ---- a2/a3: row gather (kgcnn/layers/gather.py:83,228) ---------------------------------------------------
REGISTER_OP("KgcnnGatherRows")
    .Input("x: float").Input("ids: int32").Output("y: float")
    .SetShapeFn([](shape_inference::InferenceContext* c) {
      c->set_output(0, c->MakeShape({c->Dim(c->input(1), 0), c->Dim(c->input(0), 1)}));
      return OkStatus();
    });

class KgcnnGatherRowsOp : public OpKernel {
 public:
  explicit KgcnnGatherRowsOp(OpKernelConstruction* c) : OpKernel(c) {}
  void Compute(OpKernelContext* ctx) override {
    const Tensor& x = ctx->input(0);
    const Tensor& ids = ctx->input(1);
    const int64_t E = ids.dim_size(0);
    const int F = static_cast<int>(x.dim_size(1));
    Tensor* y = nullptr;
    OP_REQUIRES_OK(ctx, ctx->allocate_output(0, TensorShape({E, F}), &y));
    KGCNN_OK_OR_RETURN(ctx, kgcnn_gather_rows(x.flat<float>().data(), x.dim_size(0), F, ids.flat<int32>().data(), 1,
                                               E, y->flat<float>().data(), F, StreamOf(ctx)));
  }
};
REGISTER_KERNEL_BUILDER(Name("KgcnnGatherRows").Device(DEVICE_GPU), KgcnnGatherRowsOp);

// ---- a4/a13/a18: fused gather * weight -> CSR segment reduce ---------------------------------------------
// (kgcnn/layers/pooling.py:63-76 + schnet_conv.py:76-78 / gcn_conv.py:87-88)
REGISTER_OP("KgcnnGatherMulSegmentReduce")
    .Input("x: float").Input("w: float").Input("rowptr: int32").Input("perm: int32").Input("gidx: int32")
    .Attr("op: int = 0").Attr("has_weight: bool = true").Attr("has_perm: bool = false")
    .Attr("has_gather: bool = true").Attr("gidx_permuted: bool = false")
    .Output("y: float")
    .SetShapeFn([](shape_inference::InferenceContext* c) {
      shape_inference::DimensionHandle n;
      TF_RETURN_IF_ERROR(c->Subtract(c->Dim(c->input(2), 0), 1, &n));
      c->set_output(0, c->MakeShape({n, c->Dim(c->input(0), 1)}));
      return OkStatus();
    });

class KgcnnGatherMulSegmentReduceOp : public OpKernel {
 public:
  explicit KgcnnGatherMulSegmentReduceOp(OpKernelConstruction* c) : OpKernel(c) {
    OP_REQUIRES_OK(c, c->GetAttr("op", &op_));
    OP_REQUIRES_OK(c, c->GetAttr("has_weight", &has_w_));
    OP_REQUIRES_OK(c, c->GetAttr("has_perm", &has_perm_));
    OP_REQUIRES_OK(c, c->GetAttr("has_gather", &has_gather_));
    OP_REQUIRES_OK(c, c->GetAttr("gidx_permuted", &gperm_));
  }
  void Compute(OpKernelContext* ctx) override {
    const Tensor& x = ctx->input(0);
    const Tensor& w = ctx->input(1);
    const Tensor& rowptr = ctx->input(2);
    const Tensor& perm = ctx->input(3);
    const Tensor& gidx = ctx->input(4);
    const int F = static_cast<int>(x.dim_size(1));
    const int64_t n_out = rowptr.dim_size(0) - 1;
    const int64_t n_el = has_gather_ ? gidx.dim_size(0) : x.dim_size(0);
    Tensor* y = nullptr;
    OP_REQUIRES_OK(ctx, ctx->allocate_output(0, TensorShape({n_out, F}), &y));
    KGCNN_OK_OR_RETURN(ctx, kgcnn_gather_mul_segment_reduce(
        x.flat<float>().data(), F, has_w_ ? w.flat<float>().data() : nullptr,
        has_w_ ? static_cast<int>(w.dim_size(1)) : 0, rowptr.flat<int32>().data(),
        has_perm_ ? perm.flat<int32>().data() : nullptr, has_gather_ ? gidx.flat<int32>().data() : nullptr, 1,
        gperm_ ? 1 : 0, n_out, n_el, op_, y->flat<float>().data(), nullptr, nullptr, 0, 0, StreamOf(ctx)));
  }
 private:
  int op_;
  bool has_w_, has_perm_, has_gather_, gperm_;
};
REGISTER_KERNEL_BUILDER(Name("KgcnnGatherMulSegmentReduce").Device(DEVICE_GPU), KgcnnGatherMulSegmentReduceOp);

// ---- VJP wrt w: two-sided gather product ----------------------------------------------------------------------
REGISTER_OP("KgcnnGather2Mul")
    .Input("a: float").Input("ia: int32").Input("b: float").Input("ib: int32").Attr("reduce_f: bool = false")
    .Output("z: float")
    .SetShapeFn([](shape_inference::InferenceContext* c) {
      c->set_output(0, c->MakeShape({c->Dim(c->input(1), 0), c->UnknownDim()}));
      return OkStatus();
    });

class KgcnnGather2MulOp : public OpKernel {
 public:
  explicit KgcnnGather2MulOp(OpKernelConstruction* c) : OpKernel(c) { OP_REQUIRES_OK(c, c->GetAttr("reduce_f", &red_)); }
  void Compute(OpKernelContext* ctx) override {
    const Tensor& a = ctx->input(0);
    const Tensor& ia = ctx->input(1);
    const Tensor& b = ctx->input(2);
    const Tensor& ib = ctx->input(3);
    const int64_t E = ia.dim_size(0);
    const int F = static_cast<int>(a.dim_size(1));
    Tensor* z = nullptr;
    OP_REQUIRES_OK(ctx, ctx->allocate_output(0, TensorShape({E, red_ ? 1 : F}), &z));
    KGCNN_OK_OR_RETURN(ctx, kgcnn_gather2_mul(a.flat<float>().data(), ia.flat<int32>().data(), 1,
                                               b.flat<float>().data(), ib.flat<int32>().data(), 1, F, E, red_ ? 1 : 0,
                                               z->flat<float>().data(), StreamOf(ctx)));
  }
 private:
  bool red_;
};
REGISTER_KERNEL_BUILDER(Name("KgcnnGather2Mul").Device(DEVICE_GPU), KgcnnGather2MulOp);
