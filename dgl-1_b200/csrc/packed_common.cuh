// Shared by packed.cu / packed_graph.cu: a minimal mirror of the C types of the reference's
// include/dgl/runtime/c_runtime_api.h and dlpack.h (ABI, not code) + the libdgl entry points resolved
// with dlsym at registration time.
#pragma once
#include <dlfcn.h>

#include <cstring>
#include <string>

#include "common.cuh"

namespace dglb200 {
namespace packed {

// --- minimal mirror of the C types of c_runtime_api.h / dlpack.h (ABI, not code) ---
typedef union {
  int64_t v_int64;
  double v_float64;
  void* v_handle;
  const char* v_str;
} RefValue;  // DGLValue (c_runtime_api.h:122-129); the two struct members are 8 bytes as well

enum { kRefInt = 0, kRefFloat = 2, kRefHandle = 3, kRefNull = 4, kRefArrayHandle = 7,
       kRefFuncHandle = 10, kRefNDArrayContainer = 13 };  // c_runtime_api.h:66-91

struct RefDLTensor {  // dlpack.h DLTensor
  void* data;
  int32_t device_type;
  int32_t device_id;
  int32_t ndim;
  uint8_t code;
  uint8_t bits;
  uint16_t lanes;
  int64_t* shape;
  int64_t* strides;
  uint64_t byte_offset;
};

typedef int (*PackedCFunc)(RefValue* args, int* type_codes, int num_args, void* ret, void* resource);
typedef void (*PackedCFuncFinalizer)(void* resource);

struct LibDgl {
  int (*FuncCreateFromCFunc)(PackedCFunc, void*, PackedCFuncFinalizer, void**) = nullptr;
  int (*FuncRegisterGlobal)(const char*, void*, int) = nullptr;
  int (*FuncFree)(void*) = nullptr;
  int (*CFuncSetReturn)(void*, RefValue*, int*, int) = nullptr;
  void (*APISetLastError)(const char*) = nullptr;
  int (*ArrayAlloc)(const int64_t*, int, int, int, int, int, int, RefDLTensor**) = nullptr;
  int (*ArrayFree)(RefDLTensor*) = nullptr;
  int (*FuncGetGlobal)(const char*, void**) = nullptr;                       // c_runtime_api.h:362
  int (*FuncCall)(void*, RefValue*, int*, int, RefValue*, int*) = nullptr;   // c_runtime_api.h:255-260
};
extern LibDgl g_dgl;

enum { kRefStr = 11 };   // c_runtime_api.h:66-91 (kStr)

int fail(const char* where);
int64_t as_int(const RefValue& v, int code);
void* as_ptr(const RefValue& v, int code);
int ret_int(void* ret, int64_t value);
int register_one(const char* name, PackedCFunc fn, int override_existing);
// packed_graph.cu: overrides of graph_index._CAPI_DGLGraph{GetAdj,InEdges_2,OutEdges_2,Edges,InDegrees,
// OutDegrees} (+ the mutators, to invalidate); returns the number registered or -1
int register_graph_overrides();

}  // namespace packed
}  // namespace dglb200
