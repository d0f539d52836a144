// glp_comm_host.inl -- multi-GPU host side, included by glprox.cu: mailbox export / attach (peer push or the NCCL
// baseline exchange).
// ------------------------------------------------------------------------------------------------
// multi-GPU: mailbox export / attach, and the one-process group driver
// ------------------------------------------------------------------------------------------------
extern "C" int glp_comm_export(glp_handle* h, int world, void* ipc_handle64_out, void** mailbox_dev_out) {
  if (h == nullptr) return fail(GLP_ERR_ARG, "null handle");
  if (world < 1 || world > MAX_WORLD) return fail(GLP_ERR_ARG, "world must be 1..%d (got %d)", MAX_WORLD, world);
  if (h->solving) return fail(GLP_ERR_STATE, "a solve is in progress on this handle");
  CU_TRY(cudaSetDevice(h->device));
  if (h->mailbox == nullptr) {
    const size_t slab_words = size_t(2) * h->g.l * h->g.ldm;
    h->mailbox_bytes = (size_t(MAIL_SLAB) + size_t(2) * MAX_WORLD * slab_words) * 8;
    CU_TRY(cudaMalloc(&h->mailbox, h->mailbox_bytes));  // plain cudaMalloc: IPC-exportable
    CU_TRY(cudaMemset(h->mailbox, 0, h->mailbox_bytes));
    CU_TRY(cudaDeviceSynchronize());
  }
  if (ipc_handle64_out != nullptr) {
    cudaIpcMemHandle_t ih;
    static_assert(sizeof(ih) == 64, "cudaIpcMemHandle_t is 64 bytes");
    CU_TRY(cudaIpcGetMemHandle(&ih, h->mailbox));
    memcpy(ipc_handle64_out, &ih, sizeof(ih));
  }
  if (mailbox_dev_out != nullptr) *mailbox_dev_out = h->mailbox;
  return GLP_OK;
}

extern "C" int glp_comm_attach(glp_handle* h, int rank, int world, int64_t n_total, const void* peers, int kind) {
  if (h == nullptr || peers == nullptr) return fail(GLP_ERR_ARG, "null argument");
  if (world < 1 || world > MAX_WORLD || rank < 0 || rank >= world) return fail(GLP_ERR_ARG, "bad rank/world %d/%d", rank, world);
  if (h->mailbox == nullptr) return fail(GLP_ERR_STATE, "glp_comm_attach before glp_comm_export");
  if (h->solving) return fail(GLP_ERR_STATE, "a solve is in progress on this handle");
  if (kind != 0 && kind != 1) return fail(GLP_ERR_ARG, "kind must be 0 (IPC handles) or 1 (device pointers)");
  if (n_total < h->g.n) return fail(GLP_ERR_ARG, "n_total (%lld) < local n (%d)", (long long)n_total, h->g.n);
  CU_TRY(cudaSetDevice(h->device));
  Comm c{};
  c.world = world;
  c.rank = rank;
  c.slab_words = int64_t(2) * h->g.l * h->g.ldm;
  for (int r = 0; r < world; ++r) {
    void* ptr = nullptr;
    if (r == rank) {
      ptr = h->mailbox;
    } else if (kind == 1) {
      ptr = static_cast<void* const*>(peers)[r];
      cudaPointerAttributes at{};
      if (ptr != nullptr && cudaPointerGetAttributes(&at, ptr) == cudaSuccess && at.type == cudaMemoryTypeDevice &&
          at.device != h->device) {
        cudaError_t e = cudaDeviceEnablePeerAccess(at.device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
          return fail(GLP_ERR_CUDA, "cudaDeviceEnablePeerAccess(%d) failed: %s", at.device, cudaGetErrorString(e));
        cudaGetLastError();
      }
    } else {
      cudaIpcMemHandle_t ih;
      memcpy(&ih, static_cast<const char*>(peers) + size_t(r) * sizeof(ih), sizeof(ih));
      cudaError_t e = cudaIpcOpenMemHandle(&ptr, ih, cudaIpcMemLazyEnablePeerAccess);
      if (e != cudaSuccess)
        return fail(GLP_ERR_CUDA, "cudaIpcOpenMemHandle(rank %d) failed: %s", r, cudaGetErrorString(e));
      h->ipc_mapped[r] = ptr;
    }
    if (ptr == nullptr) return fail(GLP_ERR_ARG, "mailbox of rank %d is null", r);
    c.mail[r] = static_cast<unsigned long long*>(ptr);
  }
  h->comm = c;
  h->g.world = world;
  h->g.n_total = n_total;
  return GLP_OK;
}

extern "C" int glp_comm_attach_nccl(glp_handle* h, void* nccl_comm, int rank, int world, int64_t n_total) {
  if (h == nullptr || nccl_comm == nullptr) return fail(GLP_ERR_ARG, "null argument");
  if (world < 1 || world > MAX_WORLD || rank < 0 || rank >= world) return fail(GLP_ERR_ARG, "bad rank/world %d/%d", rank, world);
  if (h->solving) return fail(GLP_ERR_STATE, "a solve is in progress on this handle");
  if (n_total < h->g.n) return fail(GLP_ERR_ARG, "n_total (%lld) < local n (%d)", (long long)n_total, h->g.n);
  CU_TRY(cudaSetDevice(h->device));
  if (h->nccl_stage == nullptr) {
    const size_t words = size_t(2) * h->g.l * h->g.ldm + 16;
    CU_TRY(cudaMalloc(reinterpret_cast<void**>(&h->nccl_stage), words * 8));
    CU_TRY(cudaMemsetAsync(h->nccl_stage, 0, words * 8, h->stream));
  }
  Comm c{};
  c.world = world;
  c.rank = rank;
  c.slab_words = int64_t(2) * h->g.l * h->g.ldm;
  h->comm = c;
  h->nccl_comm = nccl_comm;
  h->xch_nccl = true;
  h->g.world = world;
  h->g.n_total = n_total;
  return GLP_OK;
}
