// lt_gemm.cu — library GEMMs with cuBLASLt's training epilogues, for the ops PyTorch does not expose:
//
//   GELU_AUX_BIAS   d = gelu(a@b + bias), aux = a@b + bias          (MLP up-projection, forward)
//   DGELU_BGRAD     d = dgelu(aux) * (a@b), bias_grad = colsum(d)   (MLP down-projection dgrad, backward)
//   BGRAD_B         d (+)= a@b, bias_grad = sum_k b[k, :]           (weight gradient + bias gradient in one pass)
//
// so the GELU forward/backward kernels and the column-sum kernels of a GPT-2 MLP disappear into GEMM
// epilogues (the activations make one trip less through HBM each).  These are plain library GEMMs — the
// tcgen05 kernels are cuBLASLt's own (`nvjet_*`); this file only describes the problem to the library.
//
// cuBLASLt is NOT linked: the library is looked up at run time (`libcublasLt.so.12`, normally already
// mapped by PyTorch), so libdlion.so keeps loading in processes that do not have it; every entry point
// then returns DLION_E_UNSUPPORTED and the Python side keeps its unfused path.
//
// All matrices are row-major here (PyTorch's view); cuBLASLt is column-major, so the call computes
// d^T = op(b)^T op(a)^T with the operands swapped, which turns "bias along n" into cuBLASLt's
// "bias along the rows of D" and "sum of b over k" into its BGRADA.
// reference: HF GPT-2 MLP (Conv1D -> NewGELU -> Conv1D) forward/backward under autograd, reached from
// run_clm.py:588-639; upstream that is 2 GEMMs + ~8 elementwise/reduction kernels per direction.
#include <cublasLt.h>
#include <dlfcn.h>

#include <map>
#include <mutex>
#include <tuple>

#include "common.cuh"

namespace {

struct LtApi {
  void* so = nullptr;
  decltype(&cublasLtCreate) Create = nullptr;
  decltype(&cublasLtMatmul) Matmul = nullptr;
  decltype(&cublasLtMatmulDescCreate) DescCreate = nullptr;
  decltype(&cublasLtMatmulDescDestroy) DescDestroy = nullptr;
  decltype(&cublasLtMatmulDescSetAttribute) DescSet = nullptr;
  decltype(&cublasLtMatrixLayoutCreate) LayoutCreate = nullptr;
  decltype(&cublasLtMatrixLayoutDestroy) LayoutDestroy = nullptr;
  decltype(&cublasLtMatmulPreferenceCreate) PrefCreate = nullptr;
  decltype(&cublasLtMatmulPreferenceDestroy) PrefDestroy = nullptr;
  decltype(&cublasLtMatmulPreferenceSetAttribute) PrefSet = nullptr;
  decltype(&cublasLtMatmulAlgoGetHeuristic) Heuristic = nullptr;
  std::map<int, cublasLtHandle_t> handles;      // one per device, created on first use there
  bool ok = false;
};

LtApi g_lt;
std::once_flag g_lt_once;
std::mutex g_lt_mu;

template <typename F>
bool sym(void* so, const char* name, F* out) {
  *out = reinterpret_cast<F>(dlsym(so, name));
  return *out != nullptr;
}

void lt_load() {
  const char* names[] = {"libcublasLt.so.12", "libcublasLt.so.13", "libcublasLt.so"};
  for (const char* n : names) {
    g_lt.so = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (g_lt.so) break;
  }
  if (!g_lt.so) return;
  bool ok = sym(g_lt.so, "cublasLtCreate", &g_lt.Create) && sym(g_lt.so, "cublasLtMatmul", &g_lt.Matmul) &&
            sym(g_lt.so, "cublasLtMatmulDescCreate", &g_lt.DescCreate) &&
            sym(g_lt.so, "cublasLtMatmulDescDestroy", &g_lt.DescDestroy) &&
            sym(g_lt.so, "cublasLtMatmulDescSetAttribute", &g_lt.DescSet) &&
            sym(g_lt.so, "cublasLtMatrixLayoutCreate", &g_lt.LayoutCreate) &&
            sym(g_lt.so, "cublasLtMatrixLayoutDestroy", &g_lt.LayoutDestroy) &&
            sym(g_lt.so, "cublasLtMatmulPreferenceCreate", &g_lt.PrefCreate) &&
            sym(g_lt.so, "cublasLtMatmulPreferenceDestroy", &g_lt.PrefDestroy) &&
            sym(g_lt.so, "cublasLtMatmulPreferenceSetAttribute", &g_lt.PrefSet) &&
            sym(g_lt.so, "cublasLtMatmulAlgoGetHeuristic", &g_lt.Heuristic);
  g_lt.ok = ok;
}

// caller holds g_lt_mu
cublasLtHandle_t handle_for_current_device() {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
  auto it = g_lt.handles.find(dev);
  if (it != g_lt.handles.end()) return it->second;
  cublasLtHandle_t h = nullptr;
  if (g_lt.Create(&h) != CUBLAS_STATUS_SUCCESS) return nullptr;
  g_lt.handles.emplace(dev, h);
  return h;
}

cudaDataType_t cuda_type(int dtype) {
  switch (dtype) {
    case DLION_F32: return CUDA_R_32F;
    case DLION_BF16: return CUDA_R_16BF;
    default: return CUDA_R_16F;
  }
}

uint32_t alignment_of(const void* p) {
  const uintptr_t v = reinterpret_cast<uintptr_t>(p);
  uint32_t a = 256;
  while (a > 1 && (v & (a - 1))) a >>= 1;
  return a;
}

// The heuristic query costs tens of microseconds on the host; its answer depends only on the problem
// description, so it is cached per (shape, layout, epilogue, dtype, alignment class).
using AlgoKey = std::tuple<int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int, int, int, int, int, uint32_t, int>;
std::map<AlgoKey, cublasLtMatmulAlgo_t> g_algos;

}  // namespace

using namespace dlion;

extern "C" int dlion_lt_available(void) {
  std::call_once(g_lt_once, lt_load);
  return g_lt.ok ? 1 : 0;
}

extern "C" int dlion_lt_gemm(const dlion_lt_args* a, void* stream) {
  std::call_once(g_lt_once, lt_load);
  if (!g_lt.ok) {
    set_error("dlion_lt_gemm: cuBLASLt could not be loaded in this process");
    return DLION_E_UNSUPPORTED;
  }
  DLION_CHECK_ARG(a != nullptr, "dlion_lt_gemm: null arguments");
  DLION_CHECK_ARG(a->m > 0 && a->n > 0 && a->k > 0, "dlion_lt_gemm: empty problem");
  DLION_CHECK_ARG(a->a && a->b && a->d, "dlion_lt_gemm: null matrix");
  DLION_CHECK_ARG(a->dtype == DLION_F32 || a->dtype == DLION_BF16 || a->dtype == DLION_F16, "dlion_lt_gemm: bad dtype");
  cublasLtEpilogue_t epi = CUBLASLT_EPILOGUE_DEFAULT;
  switch (a->epilogue) {
    case DLION_LT_NONE: break;
    case DLION_LT_BIAS:
      DLION_CHECK_ARG(a->bias, "dlion_lt_gemm: BIAS needs bias");
      epi = CUBLASLT_EPILOGUE_BIAS;
      break;
    case DLION_LT_GELU_AUX_BIAS:
      DLION_CHECK_ARG(a->bias && a->aux && a->ld_aux >= a->n && a->ld_aux % 8 == 0,
                      "dlion_lt_gemm: GELU_AUX_BIAS needs bias, aux and ld_aux %% 8 == 0");
      epi = CUBLASLT_EPILOGUE_GELU_AUX_BIAS;
      break;
    case DLION_LT_DGELU_BGRAD:
      DLION_CHECK_ARG(a->bias_grad && a->aux && a->ld_aux >= a->n && a->ld_aux % 8 == 0,
                      "dlion_lt_gemm: DGELU_BGRAD needs bias_grad, aux and ld_aux %% 8 == 0");
      epi = CUBLASLT_EPILOGUE_DGELU_BGRAD;
      break;
    case DLION_LT_DGELU:
      DLION_CHECK_ARG(a->aux && a->ld_aux >= a->n && a->ld_aux % 8 == 0, "dlion_lt_gemm: DGELU needs aux and ld_aux %% 8 == 0");
      epi = CUBLASLT_EPILOGUE_DGELU;
      break;
    case DLION_LT_BGRAD_B:
      DLION_CHECK_ARG(a->bias_grad, "dlion_lt_gemm: BGRAD_B needs bias_grad");
      epi = CUBLASLT_EPILOGUE_BGRADA;          // our b is cuBLASLt's A after the row-major swap
      break;
    default:
      set_error("dlion_lt_gemm: unknown epilogue %d", a->epilogue);
      return DLION_E_BADARG;
  }
  const cudaDataType_t ty = cuda_type(a->dtype);
  std::lock_guard<std::mutex> lock(g_lt_mu);
  cublasLtHandle_t handle = handle_for_current_device();
  if (!handle) {
    set_error("dlion_lt_gemm: cublasLtCreate failed on the current device");
    return DLION_E_UNSUPPORTED;
  }
  cublasLtMatmulDesc_t desc = nullptr;
  cublasLtMatrixLayout_t la = nullptr, lb = nullptr, ld = nullptr;
  cublasLtMatmulPreference_t pref = nullptr;
  int rc = 0;
  auto fail = [&](const char* what, int st) {
    set_error("dlion_lt_gemm: %s failed (cublas status %d; m=%lld n=%lld k=%lld epilogue=%d)", what, st,
              (long long)a->m, (long long)a->n, (long long)a->k, a->epilogue);
    rc = DLION_E_UNSUPPORTED;
  };
  do {
    cublasStatus_t st = g_lt.DescCreate(&desc, CUBLAS_COMPUTE_32F, CUDA_R_32F);
    if (st != CUBLAS_STATUS_SUCCESS) { fail("MatmulDescCreate", st); break; }
    // row-major d[m,n] = op(a) op(b)  <=>  column-major d^T[n,m] = op(b)^T op(a)^T
    const cublasOperation_t opA = a->trans_b ? CUBLAS_OP_T : CUBLAS_OP_N;     // cuBLASLt A := our b
    const cublasOperation_t opB = a->trans_a ? CUBLAS_OP_T : CUBLAS_OP_N;     // cuBLASLt B := our a
    g_lt.DescSet(desc, CUBLASLT_MATMUL_DESC_TRANSA, &opA, sizeof(opA));
    g_lt.DescSet(desc, CUBLASLT_MATMUL_DESC_TRANSB, &opB, sizeof(opB));
    g_lt.DescSet(desc, CUBLASLT_MATMUL_DESC_EPILOGUE, &epi, sizeof(epi));
    if (a->epilogue == DLION_LT_BIAS || a->epilogue == DLION_LT_GELU_AUX_BIAS)
      g_lt.DescSet(desc, CUBLASLT_MATMUL_DESC_BIAS_POINTER, &a->bias, sizeof(void*));
    if (a->epilogue == DLION_LT_DGELU_BGRAD || a->epilogue == DLION_LT_BGRAD_B)
      g_lt.DescSet(desc, CUBLASLT_MATMUL_DESC_BIAS_POINTER, &a->bias_grad, sizeof(void*));
    if (a->epilogue == DLION_LT_GELU_AUX_BIAS || a->epilogue == DLION_LT_DGELU_BGRAD || a->epilogue == DLION_LT_DGELU) {
      g_lt.DescSet(desc, CUBLASLT_MATMUL_DESC_EPILOGUE_AUX_POINTER, &a->aux, sizeof(void*));
      g_lt.DescSet(desc, CUBLASLT_MATMUL_DESC_EPILOGUE_AUX_LD, &a->ld_aux, sizeof(int64_t));
    }
    // stored shapes (column-major view of the row-major buffers)
    const uint64_t a_rows = a->trans_b ? a->k : a->n, a_cols = a->trans_b ? a->n : a->k;   // our b
    const uint64_t b_rows = a->trans_a ? a->m : a->k, b_cols = a->trans_a ? a->k : a->m;   // our a
    if ((st = g_lt.LayoutCreate(&la, ty, a_rows, a_cols, a->ldb)) != CUBLAS_STATUS_SUCCESS) { fail("LayoutCreate(b)", st); break; }
    if ((st = g_lt.LayoutCreate(&lb, ty, b_rows, b_cols, a->lda)) != CUBLAS_STATUS_SUCCESS) { fail("LayoutCreate(a)", st); break; }
    if ((st = g_lt.LayoutCreate(&ld, ty, a->n, a->m, a->ldd)) != CUBLAS_STATUS_SUCCESS) { fail("LayoutCreate(d)", st); break; }
    uint32_t al = alignment_of(a->a);
    for (const void* p : {a->b, (const void*)a->d, a->bias, (const void*)a->bias_grad, (const void*)a->aux})
      if (p) { const uint32_t x = alignment_of(p); al = x < al ? x : al; }
    const AlgoKey key{a->m, a->n, a->k, a->lda, a->ldb, a->ldd, a->ld_aux, a->trans_a, a->trans_b, a->epilogue,
                      a->dtype, a->beta != 0.f ? 1 : 0, al, a->algo_rank};
    auto it = g_algos.find(key);
    if (it == g_algos.end()) {
      if ((st = g_lt.PrefCreate(&pref)) != CUBLAS_STATUS_SUCCESS) { fail("PreferenceCreate", st); break; }
      const uint64_t ws = a->workspace ? (uint64_t)a->workspace_bytes : 0;
      g_lt.PrefSet(pref, CUBLASLT_MATMUL_PREF_MAX_WORKSPACE_BYTES, &ws, sizeof(ws));
      g_lt.PrefSet(pref, CUBLASLT_MATMUL_PREF_MIN_ALIGNMENT_A_BYTES, &al, sizeof(al));
      g_lt.PrefSet(pref, CUBLASLT_MATMUL_PREF_MIN_ALIGNMENT_B_BYTES, &al, sizeof(al));
      g_lt.PrefSet(pref, CUBLASLT_MATMUL_PREF_MIN_ALIGNMENT_C_BYTES, &al, sizeof(al));
      g_lt.PrefSet(pref, CUBLASLT_MATMUL_PREF_MIN_ALIGNMENT_D_BYTES, &al, sizeof(al));
      constexpr int kMaxRank = 32;
      cublasLtMatmulHeuristicResult_t res[kMaxRank];
      int found = 0;
      const int want = a->algo_rank + 1;
      if (want < 1 || want > kMaxRank) { fail("algo_rank out of range", 0); break; }
      st = g_lt.Heuristic(handle, desc, la, lb, ld, ld, pref, want, res, &found);
      if (st != CUBLAS_STATUS_SUCCESS || found < want) { fail("AlgoGetHeuristic (no kernel / no such candidate)", st); break; }
      it = g_algos.emplace(key, res[want - 1].algo).first;
    }
    st = g_lt.Matmul(handle, desc, &a->alpha, a->b, la, a->a, lb, &a->beta, a->d, ld, a->d, ld, &it->second,
                     a->workspace, a->workspace ? (size_t)a->workspace_bytes : 0, (cudaStream_t)stream);
    if (st != CUBLAS_STATUS_SUCCESS) { fail("Matmul", st); break; }
  } while (false);
  if (pref) g_lt.PrefDestroy(pref);
  if (la) g_lt.LayoutDestroy(la);
  if (lb) g_lt.LayoutDestroy(lb);
  if (ld) g_lt.LayoutDestroy(ld);
  if (desc) g_lt.DescDestroy(desc);
  return rc;
}
