// capi_group.inl — third part of capi.cu: exa_group_*, N GPUs of one node driven by ONE host thread behind the C ABI
// ("one host thread drives N GPUs via streams"; what the reference's deviceIndex / deviceCount of
// programs/LaunchParams.h:31-32 were meant for and never did).  A group owns one renderer per device and, per frame,
//   image tiles (scene replicated):  every device renders its tiles straight into N peer-mapped row strips (strip k lives
//       on device k), then device k copies strip k to the caller's host frame over its own PCIe link;
//   sort-last (data partitioned):    every device renders its kd sub-volume to a float4 partial, then device k composites
//       pixel span k from ALL partials through peer pointers (render.cu:k_compositeFinish) and copies span k to the host.
// Ordering between devices is by CUDA events (cudaStreamWaitEvent works across the devices of a process); the one-process-
// per-GPU variant of the same pipelines (torch.distributed rendezvous, CUDA IPC, exa_peer_barrier) is owlexabrick_b200/multigpu.py.

struct exa_group {
  int n = 0, mode = EXA_PARTITION_IMAGE_TILES;
  std::vector<exa_renderer *> r;
  std::vector<cudaStream_t> stream;                          // one per device, owned
  std::vector<cudaEvent_t> evRendered;
  int width = 0, height = 0;
  std::vector<uint32_t *> strip;                             // tiles: strip k on device k
  std::vector<float4 *> partial, surf;                       // sort-last: partial image (+ surface record) of device k
  std::vector<uint32_t *> spanOut;                           // sort-last: RGBA8 frame buffer of device k (its span is valid)
};

static void groupFreeBuffers(exa_group *g)
{
  for (int k = 0; k < g->n; k++) {
    cudaSetDevice(g->r[k]->device);
    if (k < (int)g->strip.size()) cudaFree(g->strip[k]);
    if (k < (int)g->partial.size()) cudaFree(g->partial[k]);
    if (k < (int)g->surf.size()) cudaFree(g->surf[k]);
    if (k < (int)g->spanOut.size()) cudaFree(g->spanOut[k]);
  }
  g->strip.clear(); g->partial.clear(); g->surf.clear(); g->spanOut.clear();
}

extern "C" {

int exa_group_create(const exa_bricks_t *bricks, const float *const *fields, const size_t *fieldCounts, int numFields,
                     const int *cudaDevices, int numDevices, int mode, exa_group_t **out)
{
  if (!out || !cudaDevices || numDevices < 1 || numDevices > EXA_MAX_PARTS) { setError("exa_group_create: 1..16 devices"); return EXA_ERR_INVALID; }
  if (mode != EXA_PARTITION_IMAGE_TILES && mode != EXA_PARTITION_SORT_LAST) { setError("exa_group_create: mode must be image tiles or sort-last"); return EXA_ERR_INVALID; }
  exa_group *g = new exa_group; g->n = numDevices; g->mode = mode;
  for (int k = 0; k < numDevices; k++) {
    exa_renderer *r = nullptr;
    const int rc = mode == EXA_PARTITION_SORT_LAST && numDevices > 1
                     ? exa_renderer_create_partitioned(bricks, fields, fieldCounts, numFields, cudaDevices[k], k, numDevices, &r)
                     : exa_renderer_create(bricks, fields, fieldCounts, numFields, cudaDevices[k], &r);
    if (rc) { exa_group_destroy(g); return rc; }
    g->r.push_back(r);
    cudaStream_t s = nullptr; cudaEvent_t e = nullptr;
    if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess) { setError("exa_group_create: stream creation failed"); exa_group_destroy(g); return EXA_ERR_CUDA; }
    g->stream.push_back(s);                                  // owned by the group from here on (destroyed on every error path)
    if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) { setError("exa_group_create: event creation failed"); exa_group_destroy(g); return EXA_ERR_CUDA; }
    g->evRendered.push_back(e);
    r->stream = s;
    if (mode == EXA_PARTITION_IMAGE_TILES && numDevices > 1) exa_renderer_set_partition(r, EXA_PARTITION_IMAGE_TILES, k, numDevices);
  }
  // every device reads / writes the others' buffers directly (NVLink peer access)
  for (int a = 0; a < numDevices; a++)
    for (int b = 0; b < numDevices; b++) {
      if (cudaDevices[a] == cudaDevices[b]) continue;
      cudaSetDevice(cudaDevices[a]);
      int can = 0; cudaDeviceCanAccessPeer(&can, cudaDevices[a], cudaDevices[b]);
      if (!can) { setError("exa_group_create: no peer access between devices " + std::to_string(cudaDevices[a]) + " and " + std::to_string(cudaDevices[b])); exa_group_destroy(g); return EXA_ERR_CUDA; }
      const cudaError_t e = cudaDeviceEnablePeerAccess(cudaDevices[b], 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { setError(std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e)); exa_group_destroy(g); return EXA_ERR_CUDA; }
      cudaGetLastError();
    }
  *out = g;
  return EXA_OK;
}

void exa_group_destroy(exa_group_t *g)
{
  if (!g) return;
  for (int k = 0; k < (int)g->r.size(); k++) { cudaSetDevice(g->r[k]->device); if (k < (int)g->stream.size()) cudaStreamSynchronize(g->stream[k]); }
  groupFreeBuffers(g);
  for (int k = 0; k < (int)g->r.size(); k++) {
    cudaSetDevice(g->r[k]->device);
    g->r[k]->stream = nullptr;
    if (k < (int)g->evRendered.size()) cudaEventDestroy(g->evRendered[k]);
    exa_renderer_destroy(g->r[k]);
    if (k < (int)g->stream.size()) cudaStreamDestroy(g->stream[k]);
  }
  delete g;
}

int exa_group_size(const exa_group_t *g) { return g ? g->n : 0; }
exa_renderer_t *exa_group_renderer(exa_group_t *g, int i) { return (g && i >= 0 && i < g->n) ? g->r[i] : nullptr; }

int exa_group_resize_framebuffer(exa_group_t *g, int width, int height)
{
  if (!g || width <= 0 || height <= 0) { setError("exa_group_resize_framebuffer: bad argument"); return EXA_ERR_INVALID; }
  groupFreeBuffers(g);
  g->width = width; g->height = height;
  const size_t npix = (size_t)width * height;
  for (int k = 0; k < g->n; k++) { const int rc = exa_renderer_resize_framebuffer(g->r[k], nullptr, width, height); if (rc) return rc; }
  if (g->mode == EXA_PARTITION_IMAGE_TILES) {
    g->strip.assign(g->n, nullptr);
    for (int k = 0; k < g->n; k++) {
      int y0, y1; exa_strip_rows(k, g->n, height, &y0, &y1);
      CKR(cudaSetDevice(g->r[k]->device));
      CKR(cudaMalloc(&g->strip[k], std::max<size_t>((size_t)(y1 - y0) * width, 1) * 4));
    }
    for (int k = 0; k < g->n; k++) { const int rc = exa_renderer_set_color_strips(g->r[k], (void *const *)g->strip.data(), g->n); if (rc) return rc; }
  } else {
    g->partial.assign(g->n, nullptr); g->surf.assign(g->n, nullptr); g->spanOut.assign(g->n, nullptr);
    for (int k = 0; k < g->n; k++) {
      CKR(cudaSetDevice(g->r[k]->device));
      CKR(cudaMalloc(&g->partial[k], npix * 16)); CKR(cudaMalloc(&g->surf[k], npix * 16)); CKR(cudaMalloc(&g->spanOut[k], npix * 4));
    }
  }
  return EXA_OK;
}

// the per-frame setters a viewer calls every frame, forwarded to every device; everything else (iso values, clip box, ...)
// goes through exa_group_renderer(g, i) with the ordinary exa_renderer_* calls, once per device
int exa_group_update_camera(exa_group_t *g, const float pos[3], const float dir00[3], const float dirDu[3], const float dirDv[3])
{ if (!g) { setError("null group"); return EXA_ERR_INVALID; } for (exa_renderer *r : g->r) { const int rc = exa_renderer_update_camera(r, pos, dir00, dirDu, dirDv); if (rc) return rc; } return EXA_OK; }
int exa_group_update_xf(exa_group_t *g, int channel, const float *opacities128, const float *rgb128x3, float lo, float hi, float opacityScale)
{ if (!g) { setError("null group"); return EXA_ERR_INVALID; } for (exa_renderer *r : g->r) { const int rc = exa_renderer_update_xf(r, channel, opacities128, rgb128x3, lo, hi, opacityScale); if (rc) return rc; } return EXA_OK; }
int exa_group_update_frame_id(exa_group_t *g, int frameID)
{ if (!g) { setError("null group"); return EXA_ERR_INVALID; } for (exa_renderer *r : g->r) exa_renderer_update_frame_id(r, frameID); return EXA_OK; }
int exa_group_set_voxel_space_transform(exa_group_t *g, const float xfm12[12])
{ if (!g) { setError("null group"); return EXA_ERR_INVALID; } for (exa_renderer *r : g->r) { const int rc = exa_renderer_set_voxel_space_transform(r, xfm12); if (rc) return rc; } return EXA_OK; }

// Image tiles dealt by measured cost (what multigpu.TilePipeline.rebalance does across processes): `rounds` frames are
// rendered with per-tile cycle counters on, every tile's cost comes from the device that rendered it, the deterministic
// LPT deal (exa_balance_tiles) gives every device its list.  A viewer calls this every few dozen frames (cost: the frames
// plus N small copies); maxOverMeanLoad (may be NULL) receives the balance of the new deal under the measured costs.
int exa_group_rebalance(exa_group_t *g, int rounds, double *maxOverMeanLoad)
{
  if (!g || g->width <= 0) { setError("exa_group_rebalance: resize the frame buffer first"); return EXA_ERR_INVALID; }
  if (maxOverMeanLoad) *maxOverMeanLoad = 1.0;
  if (g->mode != EXA_PARTITION_IMAGE_TILES || g->n < 2) return EXA_OK;
  const int N = g->n, nt = ((g->width + 15) / 16) * ((g->height + 7) / 8);
  std::vector<uint32_t> costs(nt), one(nt), mine(nt);
  for (int it = 0; it < std::max(rounds, 1); it++) {
    for (int k = 0; k < N; k++) { const int rc = exa_renderer_enable_tile_costs(g->r[k], 1); if (rc) return rc; }
    for (int k = 0; k < N; k++) { const int rc = doRender(g->r[k], nullptr, nullptr, nullptr, false); if (rc) return rc; }
    std::fill(costs.begin(), costs.end(), 0u);
    for (int k = 0; k < N; k++) {
      int rc = exa_renderer_get_tile_costs(g->r[k], one.data(), nt); if (rc) return rc;      // synchronises device k
      for (int i = 0; i < nt; i++) costs[i] = std::max(costs[i], one[i]);                       // each tile was rendered by exactly one device
      rc = exa_renderer_enable_tile_costs(g->r[k], 0); if (rc) return rc;
    }
    double total = 0.0, worst = 0.0;
    for (int k = 0; k < N; k++) {
      int n = 0;
      int rc = exa_balance_tiles(costs.data(), nt, N, k, mine.data(), &n); if (rc) return rc;
      rc = exa_renderer_set_tile_list(g->r[k], mine.data(), n); if (rc) return rc;
      double load = 0.0; for (int i = 0; i < n; i++) load += (double)costs[mine[i]];
      total += load; worst = std::max(worst, load);
    }
    if (maxOverMeanLoad && total > 0.0) *maxOverMeanLoad = worst / (total / N);
  }
  return EXA_OK;
}

int exa_group_render(exa_group_t *g, uint32_t *hostRgba8)
{
  if (!g || !hostRgba8 || g->width <= 0) { setError("exa_group_render: resize the frame buffer first"); return EXA_ERR_INVALID; }
  const int W = g->width, H = g->height, N = g->n;
  const size_t npix = (size_t)W * H;
  // 1. every device renders its share (asynchronously, one stream per device)
  for (int k = 0; k < N; k++) {
    exa_renderer *r = g->r[k];
    int rc;
    if (g->mode == EXA_PARTITION_IMAGE_TILES) rc = doRender(r, nullptr, nullptr, nullptr, false);
    else rc = doRender(r, g->partial[k], g->surf[k], nullptr, false);
    if (rc) return rc;
    CKR(cudaEventRecord(g->evRendered[k], g->stream[k]));
  }
  // 2. device k waits for ALL devices, then finishes and reads back its part over its own PCIe link
  int order[EXA_MAX_PARTS], parts = N;
  if (g->mode == EXA_PARTITION_SORT_LAST && N > 1) { const int rc = exa_renderer_partition_order(g->r[0], order, &parts); if (rc) return rc; }
  for (int k = 0; k < N; k++) {
    exa_renderer *r = g->r[k];
    CKR(cudaSetDevice(r->device));
    for (int j = 0; j < N; j++) if (j != k) CKR(cudaStreamWaitEvent(g->stream[k], g->evRendered[j], 0));
    if (g->mode == EXA_PARTITION_IMAGE_TILES) {
      int y0, y1; exa_strip_rows(k, N, H, &y0, &y1);
      if (y1 > y0) CKR(cudaMemcpyAsync(hostRgba8 + (size_t)y0 * W, g->strip[k], (size_t)(y1 - y0) * W * 4, cudaMemcpyDeviceToHost, g->stream[k]));
    } else {
      const size_t lo = npix * k / N, hi = npix * (k + 1) / N;
      const float *pp[EXA_MAX_PARTS], *ss[EXA_MAX_PARTS];
      const int n = N > 1 ? parts : 1;
      for (int j = 0; j < n; j++) { const int src = N > 1 ? order[j] : 0; pp[j] = (const float *)g->partial[src]; ss[j] = (const float *)g->surf[src]; }
      RenderParams P; fillParams(r, P);
      const int rc = exa_renderer_composite_partials(r, pp, P.anySurf ? ss : nullptr, n, lo, hi - lo, g->spanOut[k]);
      if (rc) return rc;
      if (hi > lo) CKR(cudaMemcpyAsync(hostRgba8 + lo, g->spanOut[k] + lo, (hi - lo) * 4, cudaMemcpyDeviceToHost, g->stream[k]));
    }
  }
  for (int k = 0; k < N; k++) { CKR(cudaSetDevice(g->r[k]->device)); CKR(cudaStreamSynchronize(g->stream[k])); }
  return EXA_OK;
}

} // extern "C"
