// Host helpers: TMA tensor-map construction through the driver entry point (resolved at run time).
#include "umma.cuh"

namespace hymeta {

PFN_encodeTiled get_encode_tiled() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) ==
            cudaSuccess && qres == cudaDriverEntryPointSuccess)
      fn = (PFN_encodeTiled)p;
  }
  return fn;
}

int make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                   const int64_t* strides_elems, const uint32_t* box, bool swizzle128) {
  return make_tmap(out, base, 2, rank, dims, strides_elems, box, swizzle128);
}

int make_tmap(CUtensorMap* out, const void* base, int elem_bytes, int rank, const uint64_t* dims,
              const int64_t* strides_elems, const uint32_t* box, bool swizzle128) {
  PFN_encodeTiled enc = get_encode_tiled();
  if (!enc) return HYMETA_ERR_NO_DEVICE;
  cuuint64_t gdim[5];
  cuuint64_t gstr[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
    if (i > 0) {
      gstr[i - 1] = (cuuint64_t)strides_elems[i] * elem_bytes;
      if (gstr[i - 1] % 16) return HYMETA_ERR_INVALID_ARG;
    }
  }
  if (((uintptr_t)base) % 16) return HYMETA_ERR_INVALID_ARG;
  CUresult r = enc(out, elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base),
                   gdim, gstr, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    g_last_cuda_error = (int)r;
    return HYMETA_ERR_CUDA;
  }
  return HYMETA_OK;
}

}  // namespace hymeta
