// Probe (GPU box): do CUDA conditional graph nodes do what a device-resident LM loop needs on this driver?
//   1. WHILE node whose body is captured from a stream, condition set on the device (cudaGraphSetConditional)
//   2. a WHILE node NESTED inside the body of another (LM loop around the PCG loop)
//   3. a cooperative launch captured inside a conditional body (the dense Cholesky kernel)
// nvcc -O2 -gencode arch=compute_100a,code=sm_100a -o tools/graph_probe tools/graph_probe.cu
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;
__global__ void k_inc(int* c, int i) { if (threadIdx.x == 0 && blockIdx.x == 0) c[i] += 1; }
__global__ void k_zero(int* c, int i) { if (threadIdx.x == 0) c[i] = 0; }
__global__ void k_cond(cudaGraphConditionalHandle h, const int* c, int i, int limit) {
  if (threadIdx.x == 0) cudaGraphSetConditional(h, c[i] < limit ? 1 : 0);
}
__global__ void k_coop(int* c, int i) { cg::this_grid().sync(); if (threadIdx.x == 0 && blockIdx.x == 0) c[i] += 1; }
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("\"error\": \"%s line %d: %s\"}\n", #x, __LINE__, cudaGetErrorString(e)); return 1; } } while (0)
static int add_while(cudaGraph_t g, cudaGraphConditionalHandle* h, cudaGraph_t* body) {
  CK(cudaGraphConditionalHandleCreate(h, g, 1, cudaGraphCondAssignDefault));
  cudaGraphNodeParams np = {};
  np.type = cudaGraphNodeTypeConditional; np.conditional.handle = *h; np.conditional.type = cudaGraphCondTypeWhile; np.conditional.size = 1;
  cudaGraphNode_t node;
  CK(cudaGraphAddNode(&node, g, nullptr, 0, &np));
  *body = np.conditional.phGraph_out[0];
  return 0;
}
int main() {
  int* c; CK(cudaMalloc(&c, 64)); CK(cudaMemset(c, 0, 64));
  cudaStream_t s; CK(cudaStreamCreate(&s));
  int v[4] = {0, 0, 0, 0};
  printf("{");
  {  // 1 + 2: outer WHILE (3 rounds), inner WHILE (5 rounds each)
    cudaGraph_t g, bo, bi; cudaGraphConditionalHandle ho, hi;
    CK(cudaGraphCreate(&g, 0));
    if (add_while(g, &ho, &bo)) return 1;
    // inner node must be created in the outer body graph
    if (add_while(bo, &hi, &bi)) return 1;
    CK(cudaStreamBeginCaptureToGraph(s, bi, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed));
    k_inc<<<1, 32, 0, s>>>(c, 1); k_inc<<<1, 32, 0, s>>>(c, 2); k_cond<<<1, 32, 0, s>>>(hi, c, 1, 5);
    CK(cudaStreamEndCapture(s, nullptr));
    // outer body: [inner while] -> then reset inner counter, count the round, set outer condition.  Nodes captured into `bo` after the
    // conditional node need a dependency on it: capture with the inner node as the dependency
    cudaGraphNode_t nodes[8]; size_t nn = 8;
    CK(cudaGraphGetNodes(bo, nodes, &nn));
    CK(cudaStreamBeginCaptureToGraph(s, bo, nodes, nullptr, nn, cudaStreamCaptureModeRelaxed));
    k_zero<<<1, 32, 0, s>>>(c, 1); k_inc<<<1, 32, 0, s>>>(c, 0); k_cond<<<1, 32, 0, s>>>(ho, c, 0, 3);
    CK(cudaStreamEndCapture(s, nullptr));
    cudaGraphExec_t ge; CK(cudaGraphInstantiate(&ge, g, 0));
    CK(cudaGraphLaunch(ge, s)); CK(cudaStreamSynchronize(s));
    CK(cudaMemcpy(v, c, 16, cudaMemcpyDeviceToHost));
    printf("\"nested_while_outer\": %d, \"nested_while_inner_total\": %d, \"expect\": [3, 15], ", v[0], v[2]);
  }
  {  // 3: cooperative launch inside a conditional body
    CK(cudaMemset(c, 0, 64));
    cudaGraph_t g, b; cudaGraphConditionalHandle h;
    CK(cudaGraphCreate(&g, 0));
    if (add_while(g, &h, &b)) return 1;
    CK(cudaStreamBeginCaptureToGraph(s, b, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed));
    int i0 = 0; void* args[] = {&c, &i0};
    cudaError_t e = cudaLaunchCooperativeKernel((const void*)k_coop, dim3(64), dim3(128), args, 0, s);
    k_cond<<<1, 32, 0, s>>>(h, c, 0, 4);
    cudaError_t e2 = cudaStreamEndCapture(s, nullptr);
    cudaGraphExec_t ge; cudaError_t e3 = (e == cudaSuccess && e2 == cudaSuccess) ? cudaGraphInstantiate(&ge, g, 0) : cudaErrorUnknown;
    if (e3 == cudaSuccess) { e3 = cudaGraphLaunch(ge, s); if (e3 == cudaSuccess) e3 = cudaStreamSynchronize(s); }
    cudaMemcpy(v, c, 16, cudaMemcpyDeviceToHost);
    printf("\"coop_in_body\": \"%s / %s / %s\", \"coop_rounds\": %d", cudaGetErrorString(e), cudaGetErrorString(e2), cudaGetErrorString(e3), v[0]);
    cudaGetLastError();
  }
  printf("}\n");
  return 0;
}
