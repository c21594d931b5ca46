// graph_index._CAPI_DGLGraphUniformSampling{,2,4,...,128} (src/graph/graph_apis.cc:434-472) behind the
// reference's own names: (handle, seeds x k, neigh_type, num_hops, num_neighbors, num_valid_seeds) -> an indexed
// getter over 5k entries: [0,k) subgraph handles, [k,2k) induced vertices, [2k,3k) induced edges, [3k,4k) layer
// ids, [4k,5k) sample probabilities (ConvertSubgraphToPackedFunc, :71-92) - what NSSubgraphLoader._prefetch
// (python/dgl/contrib/sampling/sampler.py:45-64) consumes, so the reference's NeighborSampler needs no change.
// The sampling itself is dglb200_sample_subgraph over a (row, col)-sorted CSR of the handle's edges kept on the
// GPU; the subgraph HANDLE has to be one of the reference's own graph objects, and the only constructor its C
// ABI offers is _CAPI_DGLGraphCreate from an edge list (:105-119) - two std::sorts on the host per subgraph,
// which the reference's sampler avoids by wrapping its CSR directly.  Included by packed_graph.cu.
#pragma once
#include <atomic>
#include <ctime>

namespace dglb200 {
namespace packed {
namespace {

struct SampleIndex {            // per read-only graph handle and neighbour type
  int64_t n = 0, e = 0;
  DevCsr csr;                   // ImmutableGraph::CSR::FromEdges order (immutable_graph.cc:206-237)
  DevMem layer, bitmap, word_prefix;
};
std::unordered_map<void*, SampleIndex*> g_sample_in, g_sample_out;
struct SampleOriginals {
  void *create = nullptr, *is_multigraph = nullptr, *is_readonly = nullptr;
} g_sorig;
std::atomic<uint64_t> g_sample_calls{0};

void drop_sample_index(void* handle) {
  std::lock_guard<std::mutex> lock(g_mu);
  for (auto* m : {&g_sample_in, &g_sample_out}) {
    auto it = m->find(handle);
    if (it != m->end()) {
      delete it->second;
      m->erase(it);
    }
  }
}

// Edges of the handle as (src, dst, eid) CPU arrays through the reference's original Edges("") - for an
// immutable graph that is out-CSR order WITH its edge ids (immutable_graph.cc:467-475).
SampleIndex* get_sample_index(void* handle, bool in_type) {
  {
    std::lock_guard<std::mutex> lock(g_mu);
    auto& m = in_type ? g_sample_in : g_sample_out;
    auto it = m.find(handle);
    if (it != m.end()) return it->second;
  }
  const int64_t n = call_count(g_orig.num_vertices, handle), e = call_count(g_orig.num_edges, handle);
  if (n <= 0 || e < 0) {
    set_error("cannot query the graph handle (NumVertices / NumEdges failed)");
    return nullptr;
  }
  RefValue order, ret;
  order.v_str = "";
  int rcode = kRefNull;
  if (call1(g_orig.edges, handle, &order, kRefStr, &ret, &rcode) != 0 || rcode != kRefFuncHandle) {
    set_error("the reference's _CAPI_DGLGraphEdges failed");
    return nullptr;
  }
  void* getter = ret.v_handle;
  RefDLTensor* arr[3] = {nullptr, nullptr, nullptr};
  bool ok = true;
  for (int i = 0; i < 3 && ok; ++i) {
    RefValue a, r;
    int c = kRefInt, rc = kRefNull;
    a.v_int64 = i;
    ok = g_dgl.FuncCall(getter, &a, &c, 1, &r, &rc) == 0 && rc == kRefNDArrayContainer;
    if (ok) arr[i] = static_cast<RefDLTensor*>(r.v_handle);
  }
  g_dgl.FuncFree(getter);
  SampleIndex* sx = new SampleIndex();
  sx->n = n;
  sx->e = e;
  DevMem d[3], ws;
  size_t ws_bytes = 0;
  const int64_t words = (n + 31) / 32;
  for (int i = 0; i < 3 && ok; ++i) {
    ok = d[i].alloc((size_t)e * 8) == 0 &&
         (e == 0 || cudaMemcpy(d[i].p, (char*)arr[i]->data + arr[i]->byte_offset, (size_t)e * 8,
                               cudaMemcpyHostToDevice) == cudaSuccess);
  }
  for (auto* t : arr)
    if (t) g_dgl.ArrayFree(t);
  if (ok) ok = dglb200_csr_build_workspace_bytes(n, e, &ws_bytes) == 0 && ws.alloc(ws_bytes) == 0 &&
               sx->csr.ptr.alloc((size_t)(n + 1) * 8) == 0 && sx->csr.idx.alloc((size_t)e * 4) == 0 &&
               sx->csr.eid.alloc((size_t)e * 4) == 0 && sx->layer.alloc((size_t)n * 4) == 0 &&
               sx->bitmap.alloc((size_t)words * 4) == 0 && sx->word_prefix.alloc((size_t)(words + 1) * 4) == 0;
  if (ok) {
    const int64_t* row = (in_type ? d[1] : d[0]).as<int64_t>();
    const int64_t* col = (in_type ? d[0] : d[1]).as<int64_t>();
    ok = dglb200_csr_build(n, n, e, row, col, d[2].as<int64_t>(), 1, sx->csr.ptr.as<int64_t>(),
                           sx->csr.idx.as<int32_t>(), sx->csr.eid.as<int32_t>(), ws.p, ws_bytes, nullptr) == 0 &&
         cudaMemset(sx->layer.p, 0xFF, (size_t)n * 4) == cudaSuccess &&
         cudaMemset(sx->bitmap.p, 0, (size_t)words * 4) == cudaSuccess;
  }
  if (!ok) {
    if (dglb200_last_error()[0] == '\0') set_error("building the device sampling index failed");
    delete sx;
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(g_mu);
  (in_type ? g_sample_in : g_sample_out)[handle] = sx;
  return sx;
}

// ---- result object: handles + arrays behind one getter --------------------------------------------------
struct SampleResult {
  int k = 0;
  std::vector<void*> graphs;                       // reference graph handles (owned by the Python side once read)
  std::vector<RefDLTensor*> nodes, edges, layers, probs;
};
void sample_result_free(void* p) {
  SampleResult* r = static_cast<SampleResult*>(p);
  for (auto* v : {&r->nodes, &r->edges, &r->layers, &r->probs})
    for (auto* t : *v)
      if (t) g_dgl.ArrayFree(t);
  delete r;
}
int pf_sample_get(RefValue* a, int* c, int n, void* ret, void* resource) {
  SampleResult* r = static_cast<SampleResult*>(resource);
  const int64_t which = n >= 1 ? as_int(a[0], c[0]) : -1;
  const int64_t k = r->k;
  if (which < 0 || which >= 5 * k) {
    set_error("invalid choice");
    return fail("sampling getter");
  }
  RefValue v;
  int code;
  if (which < k) {
    v.v_handle = r->graphs[which];
    code = kRefHandle;
  } else {
    const std::vector<RefDLTensor*>* lists[4] = {&r->nodes, &r->edges, &r->layers, &r->probs};
    v.v_handle = (*lists[which / k - 1])[which % k];
    code = kRefNDArrayContainer;
  }
  return g_dgl.CFuncSetReturn(ret, &v, &code, 1);
}

RefDLTensor* cpu_f32_zeros(int64_t len) {
  RefDLTensor* t = nullptr;
  if (g_dgl.ArrayAlloc(&len, 1, 2, 32, 1, 1, 0, &t) != 0) return nullptr;
  if (len > 0) memset(t->data, 0, (size_t)len * 4);
  return t;
}

// one seed array -> one subgraph; appends to `res`
int sample_one(void* handle, SampleIndex* sx, const RefDLTensor* seeds, bool in_type, int hops, int64_t fanout,
               bool multigraph, SampleResult* res) {
  const int64_t k = seeds->shape[0];
  if (k == 0) {                                    // padding entry (graph_index.py:1021-1024)
    res->graphs.push_back(nullptr);
    res->nodes.push_back(cpu_ids(0));
    res->edges.push_back(cpu_ids(0));
    res->layers.push_back(cpu_ids(0));
    res->probs.push_back(cpu_f32_zeros(0));
    return 0;
  }
  DevMem d_seeds, ws, wide;
  if (d_seeds.alloc((size_t)k * 8) != 0 ||
      cudaMemcpy(d_seeds.p, (char*)seeds->data + seeds->byte_offset, (size_t)k * 8, cudaMemcpyHostToDevice) != cudaSuccess) {
    set_error("cannot copy the seeds to the device");
    return -1;
  }
  const uint64_t seed = (uint64_t)time(nullptr) * 0x9E3779B97F4A7C15ull + g_sample_calls.fetch_add(1);
  dglb200_sampled_subgraph_t out;
  size_t bytes = 1u << 20;
  int rc = -1;
  for (int attempt = 0; attempt < 2 * hops + 4; ++attempt) {
    if (ws.alloc(bytes) != 0) { set_error("cudaMalloc failed"); return -1; }
    rc = dglb200_sample_subgraph(sx->n, sx->csr.ptr.as<int64_t>(), sx->csr.idx.as<int32_t>(), sx->csr.eid.as<int32_t>(),
                                 k, d_seeds.as<int64_t>(), fanout, hops, seed, sx->layer.as<int32_t>(),
                                 sx->bitmap.as<uint32_t>(), sx->word_prefix.as<int32_t>(), ws.p, bytes, &out, nullptr);
    if (rc == 0 || out.needed_bytes <= bytes) break;
    bytes = out.needed_bytes;
  }
  if (rc != 0) return -1;
  const int64_t ns = out.n_nodes, ne = out.n_edges;
  RefDLTensor* nodes = from_device(out.nodes, ns);
  RefDLTensor* layers = from_device(out.layers, ns);
  RefDLTensor* edges = from_device(out.induced_edges, ne);
  RefDLTensor* rows = from_device(out.sub_rows, ne);
  RefDLTensor* cols = nullptr;
  if (wide.alloc((size_t)(ne > 0 ? ne : 1) * 8) == 0 &&
      dglb200_widen_i32_to_i64(ne, out.sub_indices, wide.as<int64_t>(), nullptr) == 0)
    cols = from_device(wide.as<int64_t>(), ne);
  RefDLTensor* ids = iota(ne);
  RefDLTensor* probs = cpu_f32_zeros(ns);
  void* sub_handle = nullptr;
  bool ok = nodes && layers && edges && rows && cols && ids && probs;
  if (ok) {
    // the subgraph as one of the reference's own read-only graphs: _CAPI_DGLGraphCreate(src, dst, eid, multigraph,
    // num_nodes, readonly) (graph_apis.cc:105-119); "in": the sampled row is the destination
    RefValue a[6], r;
    int c[6] = {kRefArrayHandle, kRefArrayHandle, kRefArrayHandle, kRefInt, kRefInt, kRefInt}, rcode = kRefNull;
    a[0].v_handle = in_type ? cols : rows;
    a[1].v_handle = in_type ? rows : cols;
    a[2].v_handle = ids;
    a[3].v_int64 = multigraph ? 1 : 0;
    a[4].v_int64 = ns;
    a[5].v_int64 = 1;
    ok = g_dgl.FuncCall(g_sorig.create, a, c, 6, &r, &rcode) == 0 && rcode == kRefHandle;
    if (ok) sub_handle = r.v_handle;
    else set_error("the reference's _CAPI_DGLGraphCreate failed for the sampled subgraph");
  }
  for (auto* t : {rows, cols, ids})
    if (t) g_dgl.ArrayFree(t);
  if (!ok) {
    for (auto* t : {nodes, layers, edges, probs})
      if (t) g_dgl.ArrayFree(t);
    if (dglb200_last_error()[0] == '\0') set_error("copying the sampled subgraph to the host failed");
    return -1;
  }
  res->graphs.push_back(sub_handle);
  res->nodes.push_back(nodes);
  res->edges.push_back(edges);
  res->layers.push_back(layers);
  res->probs.push_back(probs);
  return 0;
}

int pf_uniform_sampling(RefValue* a, int* c, int n, void* ret, void* resource) {
  const char* who = "graph_index._CAPI_DGLGraphUniformSampling[b200]";
  const int k = (int)(intptr_t)resource;
  if (n != k + 5) { set_error("expects (handle, %d seed arrays, neigh_type, num_hops, num_neighbors, num_valid_seeds)", k); return fail(who); }
  void* handle = as_ptr(a[0], c[0]);
  if (c[k + 1] != kRefStr) { set_error("neigh_type must be a string"); return fail(who); }
  const std::string ntype = a[k + 1].v_str;
  const int hops = (int)as_int(a[k + 2], c[k + 2]);
  const int64_t fanout = as_int(a[k + 3], c[k + 3]);
  const int64_t valid = as_int(a[k + 4], c[k + 4]);
  if (valid > k) { set_error("num_valid_seeds > num_seeds"); return fail(who); }       // graph_apis.cc:449
  if (call_count(g_sorig.is_readonly, handle) != 1) {
    set_error("sampling isn't implemented in mutable graph");                         // graph_apis.cc:448
    return fail(who);
  }
  const bool in_type = ntype == "in";                                                 // immutable_graph.cc:780
  SampleIndex* sx = get_sample_index(handle, in_type);
  if (sx == nullptr) return fail(who);
  const bool multigraph = call_count(g_sorig.is_multigraph, handle) == 1;
  SampleResult* res = new SampleResult();
  res->k = k;
  static const RefDLTensor kEmpty = {nullptr, 1, 0, 1, 0, 64, 1, nullptr, nullptr, 0};
  int64_t zero = 0;
  RefDLTensor empty = kEmpty;
  empty.shape = &zero;
  for (int i = 0; i < k; ++i) {
    const RefDLTensor* seeds = static_cast<const RefDLTensor*>(a[1 + i].v_handle);
    const bool is_arr = c[1 + i] == kRefArrayHandle || c[1 + i] == kRefNDArrayContainer;
    if (i < valid && (!is_arr || !valid_ids(seeds))) {
      sample_result_free(res);
      set_error("Invalid vertex id array.");
      return fail(who);
    }
    if (sample_one(handle, sx, i < valid ? seeds : &empty, in_type, hops, fanout, multigraph, res) != 0) {
      sample_result_free(res);
      return fail(who);
    }
  }
  void* getter = nullptr;
  if (g_dgl.FuncCreateFromCFunc(pf_sample_get, res, sample_result_free, &getter) != 0) {
    sample_result_free(res);
    return fail(who);
  }
  RefValue v;
  v.v_handle = getter;
  int code = kRefFuncHandle;
  const int rc = g_dgl.CFuncSetReturn(ret, &v, &code, 1);
  g_dgl.FuncFree(getter);
  return rc;
}

int register_sampling_overrides() {
  if (get_global("graph_index._CAPI_DGLGraphCreate", &g_sorig.create) != 0) return -1;
  if (get_global("graph_index._CAPI_DGLGraphIsMultigraph", &g_sorig.is_multigraph) != 0) return -1;
  if (get_global("graph_index._CAPI_DGLGraphIsReadonly", &g_sorig.is_readonly) != 0) return -1;
  int n = 0;
  for (int k = 1; k <= 128; k *= 2) {
    const std::string name = std::string("graph_index._CAPI_DGLGraphUniformSampling") + (k == 1 ? "" : std::to_string(k));
    void* h = nullptr;
    if (g_dgl.FuncCreateFromCFunc(pf_uniform_sampling, (void*)(intptr_t)k, nullptr, &h) != 0) return -1;
    const int rc = g_dgl.FuncRegisterGlobal(name.c_str(), h, 1);
    g_dgl.FuncFree(h);
    if (rc != 0) { set_error("cannot register %s", name.c_str()); return -1; }
    ++n;
  }
  return n;
}

}  // namespace
}  // namespace packed
}  // namespace dglb200
