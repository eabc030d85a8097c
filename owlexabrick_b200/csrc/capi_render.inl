// capi_render.inl — second half of capi.cu: render(), sampling, tracer, sort-last helpers.

static int refreshMasks(exa_renderer *r, RenderParams &P)
{
  if (!r->masksDirty) return EXA_OK;
  // space skipping is forced off while a contour plane is enabled (OptixRenderer.cpp:418-432)
  bool contour = false;
  for (int i = 0; i < EXA_MAX_CONTOUR_PLANES; i++) contour |= r->fs.contourPlane[i].enabled != 0;
  const uint8_t *inPart = r->partMode == EXA_PARTITION_SORT_LAST ? r->d_inPart : nullptr;
  int n = launchRegionActivity(P, r->abr.numKdNodes, r->d_kdParent, !contour && r->spaceSkipping, inPart,
                               r->d_volActive, r->d_isoActive, r->d_regionActive, r->stream);
  if (n < 0) { setError(std::string("activity kernels failed: ") + cudaGetErrorString(cudaGetLastError())); return EXA_ERR_CUDA; }
  r->launches += n;
  r->masksDirty = false;
  return EXA_OK;
}

// Streamline cones: (re)build the uniform grid the frame kernel's traceCones walks.  Runs when the
// traces changed (reset / advance), i.e. where the reference rebuilds its streamline BVH
// (OptixRenderer.cpp:539-543).  Per-trace boxes (k_traceBounds) give the grid bounds; k_coneGrid
// counts, the host turns counts into cell starts (1 MB each way), k_coneGrid fills.
static const int CONE_GRID = 64;
static int buildConeGrid(exa_renderer *r, const RenderParams &P)
{
  const size_t nc = (size_t)CONE_GRID * CONE_GRID * CONE_GRID;
  r->coneGridValid = false; r->traceBoundsDirty = false;
  if (r->numTraces > r->traceBoundsCap) {
    CKR(cudaStreamSynchronize(r->stream)); cudaFree(r->d_traceBounds); r->d_traceBounds = nullptr;
    CKR(cudaMalloc(&r->d_traceBounds, (size_t)r->numTraces * 24)); r->traceBoundsCap = r->numTraces;
  }
  int k = launchTraceBounds(P.traces, P.numTraces, P.numTimesteps, P.currentTimestep, r->d_traceBounds, r->stream);
  if (k < 0) { setError(std::string("trace bounds kernel failed: ") + cudaGetErrorString(cudaGetLastError())); return EXA_ERR_CUDA; }
  r->launches += k;
  std::vector<float> hb((size_t)r->numTraces * 6);
  CKR(cudaMemcpyAsync(hb.data(), r->d_traceBounds, hb.size() * 4, cudaMemcpyDeviceToHost, r->stream));
  CKR(cudaStreamSynchronize(r->stream));
  float lo[3] = {3e38f, 3e38f, 3e38f}, hi[3] = {-3e38f, -3e38f, -3e38f}; bool any = false;
  for (int i = 0; i < r->numTraces; i++) {
    const float *b = hb.data() + 6 * (size_t)i;
    if (!(b[0] <= b[3])) continue;
    any = true;
    for (int d = 0; d < 3; d++) { lo[d] = std::min(lo[d], b[d]); hi[d] = std::max(hi[d], b[3 + d]); }
  }
  if (!any) return EXA_OK;
  float inv[3];
  for (int d = 0; d < 3; d++) {
    const float pad = 1e-3f * (hi[d] - lo[d]) + 1e-3f;
    lo[d] -= pad; hi[d] += pad;
    r->coneLo[d] = lo[d]; r->coneCell[d] = (hi[d] - lo[d]) / CONE_GRID; inv[d] = 1.f / r->coneCell[d];
  }
  if (!r->d_coneStart) { CKR(cudaMalloc(&r->d_coneStart, (nc + 1) * sizeof(int))); CKR(cudaMalloc(&r->d_coneCursor, nc * sizeof(int))); }
  CKR(cudaMemsetAsync(r->d_coneCursor, 0, nc * sizeof(int), r->stream));
  k = launchConeGrid(P.traces, P.numTraces, P.numTimesteps, P.currentTimestep, CONE_GRID, lo, inv, r->d_coneCursor, nullptr, 0, r->stream);
  if (k < 0) { setError(std::string("cone grid kernel failed: ") + cudaGetErrorString(cudaGetLastError())); return EXA_ERR_CUDA; }
  r->launches += k;
  std::vector<int> h(nc + 1);
  CKR(cudaMemcpyAsync(h.data(), r->d_coneCursor, nc * sizeof(int), cudaMemcpyDeviceToHost, r->stream));
  CKR(cudaStreamSynchronize(r->stream));
  size_t total = 0;
  for (size_t c = 0; c < nc; c++) { const int n = h[c]; h[c] = (int)total; total += (size_t)n; }
  h[nc] = (int)total;
  if (total == 0) return EXA_OK;
  if (total > (size_t)0x7fffffff) { setError("cone grid: too many entries"); return EXA_ERR_LIMIT; }
  if (total > r->coneListCap) { cudaFree(r->d_coneList); r->d_coneList = nullptr; CKR(cudaMalloc(&r->d_coneList, total * sizeof(int))); r->coneListCap = total; }
  CKR(cudaMemcpyAsync(r->d_coneStart, h.data(), (nc + 1) * sizeof(int), cudaMemcpyHostToDevice, r->stream));
  CKR(cudaMemcpyAsync(r->d_coneCursor, h.data(), nc * sizeof(int), cudaMemcpyHostToDevice, r->stream));
  k = launchConeGrid(P.traces, P.numTraces, P.numTimesteps, P.currentTimestep, CONE_GRID, lo, inv, r->d_coneCursor, r->d_coneList, 1, r->stream);
  if (k < 0) { setError(std::string("cone grid kernel failed: ") + cudaGetErrorString(cudaGetLastError())); return EXA_ERR_CUDA; }
  r->launches += k;
  CKR(cudaStreamSynchronize(r->stream));                     // h[] is pageable and goes out of scope
  r->coneGridValid = true;
  return EXA_OK;
}

static int doRender(exa_renderer *r, float4 *partialOut, float4 *partialSurf, ExaStats *stats, bool sync)
{
  if (!r) { setError("null renderer"); return EXA_ERR_INVALID; }
  if (!r->d_accum || r->width <= 0) { setError("render: resizeFrameBuffer was never called"); return EXA_ERR_INVALID; }
  CKR(cudaSetDevice(r->device));
  RenderParams P; fillParams(r, P);
  int rc = refreshMasks(r, P); if (rc) return rc;
  P.partialOut = partialOut; P.partialSurf = P.anySurf ? partialSurf : nullptr;
  if (partialOut && P.anySurf && !partialSurf) { setError("render_partial: a surface is enabled (iso / plane / mesh / cones): pass a surface buffer too"); return EXA_ERR_INVALID; }
  if (P.traces) {                                            // uniform grid over the visible cones
    if (r->traceBoundsDirty) { rc = buildConeGrid(r, P); if (rc) return rc; }
    if (r->coneGridValid) {
      P.coneStart = r->d_coneStart; P.coneList = r->d_coneList; P.coneG = CONE_GRID;
      for (int d = 0; d < 3; d++) { P.coneLo[d] = r->coneLo[d]; P.coneCell[d] = r->coneCell[d]; P.coneInvCell[d] = 1.f / r->coneCell[d]; }
    } else P.traces = nullptr;                               // every visible segment has a dead end
  }
  if (stats && P.gradCD) { setError("render_stats: the instrumented frame is not available with central-difference gradients"); return EXA_ERR_INVALID; }
  if (stats) { CKR(cudaMemsetAsync(r->d_stats, 0, 64, r->stream)); P.stats = r->d_stats; }
  if (partialOut && sortLastParts(r) && P.anySurf != 2 && !P.tileList) {
    // Sort-last partial: only the tiles that hold the projection of this part's box are launched; the rest of the partial
    // is not written, and the composite does not read it (exa_renderer_composite_partials derives the same rectangles).
    // Replicated surfaces (meshes, planes, cones: anySurf == 2) can lie anywhere on the screen: whole frame then.
    if (P.sceneRect[2] < P.sceneRect[0] || P.sceneRect[3] < P.sceneRect[1]) return EXA_OK;          // nothing of this part is on the screen
    P.gridX0 = P.sceneRect[0] / 16; P.gridY0 = P.sceneRect[1] / 8;
    P.gridW = P.sceneRect[2] / 16 - P.gridX0 + 1; P.gridH = P.sceneRect[3] / 8 - P.gridY0 + 1;
    const int g[4] = {P.gridX0, P.gridY0, P.gridW, P.gridH};
    if (memcmp(g, r->lastGrid, sizeof g)) { r->tileOrderValid = false; memcpy(r->lastGrid, g, sizeof g); }   // other tiles than the recorded costs describe
  } else if (r->lastGrid[2]) { r->tileOrderValid = false; memset(r->lastGrid, 0, sizeof r->lastGrid); }
  if (r->d_tileOrder[0] && !P.tileList) {                    // cost-ordered tile schedule (render.cu); an explicit tile list is already ordered
    if (!r->tileOrderValid)                                  // (re)start: both buffers' counters to zero, once
      for (int i = 0; i < 2; i++) CKR(cudaMemsetAsync(r->d_tileOrder[i], 0, offsetof(TileOrder, list), r->stream));
    if (r->tileOrderValid) P.tileOrderIn = r->d_tileOrder[r->tileOrderCur];
    P.tileOrderOut = r->d_tileOrder[r->tileOrderCur ^ 1];
    r->tileOrderCur ^= 1; r->tileOrderValid = true;
  }
  const int nLaunched = launchRenderFrame(P, r->stream);
  if (nLaunched < 0) { setError(std::string("frame kernel launch failed: ") + cudaGetErrorString(cudaGetLastError())); return EXA_ERR_CUDA; }
  r->launches += nLaunched;
  if (stats) {
    unsigned long long h[4];
    CKR(cudaMemcpyAsync(h, r->d_stats, 32, cudaMemcpyDeviceToHost, r->stream));
    CKR(cudaStreamSynchronize(r->stream));
    stats->samples = h[0]; stats->brickVisits = h[1]; stats->cornerLoads = h[2]; stats->segments = h[3];
  } else if (sync) CKR(cudaStreamSynchronize(r->stream));
  return EXA_OK;
}

extern "C" {

int exa_renderer_render(exa_renderer_t *r) { return doRender(r, nullptr, nullptr, nullptr, true); }
int exa_renderer_render_async(exa_renderer_t *r) { return doRender(r, nullptr, nullptr, nullptr, false); }
int exa_renderer_render_stats(exa_renderer_t *r, ExaStats *stats)
{
  if (!stats) { setError("null stats"); return EXA_ERR_INVALID; }
  return doRender(r, nullptr, nullptr, stats, true);
}
int exa_renderer_get_tile_cost_histogram(exa_renderer_t *r, int counts[16])
{
  if (!r || !counts) { setError("null argument"); return EXA_ERR_INVALID; }
  memset(counts, 0, 16 * sizeof(int));
  if (!r->d_tileOrder[0] || !r->tileOrderValid) return EXA_OK;
  CKR(cudaSetDevice(r->device));
  CKR(cudaStreamSynchronize(r->stream));
  CKR(cudaMemcpy(counts, r->d_tileOrder[r->tileOrderCur]->count, 16 * sizeof(int), cudaMemcpyDeviceToHost));
  return EXA_OK;
}
int exa_renderer_debug_counters(exa_renderer_t *r, uint64_t out[4])
{
  if (!r || !out) { setError("null argument"); return EXA_ERR_INVALID; }
  CKR(cudaSetDevice(r->device));
  CKR(cudaStreamSynchronize(r->stream));
  CKR(cudaMemcpy(out, r->d_stats + 4, 32, cudaMemcpyDeviceToHost));
  return EXA_OK;
}
int exa_renderer_sync(exa_renderer_t *r)
{
  if (!r) { setError("null renderer"); return EXA_ERR_INVALID; }
  CKR(cudaSetDevice(r->device)); CKR(cudaStreamSynchronize(r->stream));
  if (r->copyStream) CKR(cudaStreamSynchronize(r->copyStream));
  return EXA_OK;
}

int exa_renderer_read_rgba8(exa_renderer_t *r, uint32_t *hostOut)
{
  if (!r || !hostOut || !r->d_rgba8) { setError("read_rgba8: no framebuffer"); return EXA_ERR_INVALID; }
  CKR(cudaSetDevice(r->device));
  CKR(cudaMemcpyAsync(hostOut, r->d_rgba8, (size_t)r->width * r->height * 4, cudaMemcpyDeviceToHost, r->stream));
  CKR(cudaStreamSynchronize(r->stream)); return EXA_OK;
}
int exa_renderer_read_rgba8_async(exa_renderer_t *r, uint32_t *pinnedHostOut)
{
  if (!r || !pinnedHostOut || !r->d_rgba8) { setError("read_rgba8_async: no framebuffer"); return EXA_ERR_INVALID; }
  if (!r->ownRgba8) { setError("read_rgba8_async needs a renderer-owned frame buffer (resizeFrameBuffer(NULL, ..))"); return EXA_ERR_INVALID; }
  CKR(cudaSetDevice(r->device));
  const size_t bytes = (size_t)r->width * r->height * 4;
  if (!r->copyStream) {
    CKR(cudaStreamCreateWithFlags(&r->copyStream, cudaStreamNonBlocking));
    CKR(cudaEventCreateWithFlags(&r->evRendered, cudaEventDisableTiming));
    for (int i = 0; i < 2; i++) CKR(cudaEventCreateWithFlags(&r->evCopied[i], cudaEventDisableTiming));
  }
  if (!r->d_rgba8Alt) CKR(cudaMalloc(&r->d_rgba8Alt, bytes));
  // copy THIS frame on the copy stream, flip to the other buffer for the next frame, and make
  // the render stream wait until that buffer's previous copy has left
  CKR(cudaEventRecord(r->evRendered, r->stream));
  CKR(cudaStreamWaitEvent(r->copyStream, r->evRendered, 0));
  CKR(cudaMemcpyAsync(pinnedHostOut, r->d_rgba8, bytes, cudaMemcpyDeviceToHost, r->copyStream));
  CKR(cudaEventRecord(r->evCopied[r->rbCur], r->copyStream));
  std::swap(r->d_rgba8, r->d_rgba8Alt); r->rbCur ^= 1;
  CKR(cudaStreamWaitEvent(r->stream, r->evCopied[r->rbCur], 0));
  return EXA_OK;
}
int exa_renderer_write_accum(exa_renderer_t *r, const float *hostIn)
{
  if (!r || !hostIn || !r->d_accum) { setError("write_accum: no framebuffer"); return EXA_ERR_INVALID; }
  CKR(cudaSetDevice(r->device));
  CKR(cudaMemcpyAsync(r->d_accum, hostIn, (size_t)r->width * r->height * 16, cudaMemcpyHostToDevice, r->stream));
  CKR(cudaStreamSynchronize(r->stream)); return EXA_OK;
}
int exa_renderer_read_accum(exa_renderer_t *r, float *hostOut)
{
  if (!r || !hostOut || !r->d_accum) { setError("read_accum: no framebuffer"); return EXA_ERR_INVALID; }
  CKR(cudaSetDevice(r->device));
  CKR(cudaMemcpyAsync(hostOut, r->d_accum, (size_t)r->width * r->height * 16, cudaMemcpyDeviceToHost, r->stream));
  CKR(cudaStreamSynchronize(r->stream)); return EXA_OK;
}
int exa_renderer_device_buffers(exa_renderer_t *r, void **accum, void **rgba8)
{
  if (!r) { setError("null renderer"); return EXA_ERR_INVALID; }
  if (accum) *accum = r->d_accum;
  if (rgba8) *rgba8 = r->d_rgba8;
  return EXA_OK;
}

// ---- point sampling ------------------------------------------------------------------
static int samplePointsDev(exa_renderer *r, const float *d_xyz, size_t n, int channel, float *d_values, uint8_t *d_valid, float *d_grad, ExaStats *stats)
{
  if (!r) { setError("null renderer"); return EXA_ERR_INVALID; }
  if (channel < 0 || channel >= r->numFields) { setError("sample_points: channel out of range"); return EXA_ERR_INVALID; }
  if (n && !d_xyz) { setError("sample_points: null points"); return EXA_ERR_INVALID; }
  CKR(cudaSetDevice(r->device));
  RenderParams RP; fillParams(r, RP);
  SampleParams P; P.scene = RP.scene; P.xyz = d_xyz; P.n = n; P.channel = channel; P.values = d_values; P.valid = d_valid; P.grad = d_grad; P.stats = nullptr;
  if (stats) { CKR(cudaMemsetAsync(r->d_stats, 0, 64, r->stream)); P.stats = r->d_stats; }
  const int k = launchSamplePoints(P, r->stream);
  if (k < 0) { setError(std::string("sample kernel launch failed: ") + cudaGetErrorString(cudaGetLastError())); return EXA_ERR_CUDA; }
  r->launches += k;
  if (stats) {
    unsigned long long h[4];
    CKR(cudaMemcpyAsync(h, r->d_stats, 32, cudaMemcpyDeviceToHost, r->stream));
    CKR(cudaStreamSynchronize(r->stream));
    stats->samples = h[0]; stats->brickVisits = h[1]; stats->cornerLoads = h[2]; stats->segments = h[3];
  }
  return EXA_OK;
}

int exa_sample_points_device(exa_renderer_t *r, const float *d_xyz, size_t n, int channel, float *d_values, uint8_t *d_valid, float *d_grad)
{ return samplePointsDev(r, d_xyz, n, channel, d_values, d_valid, d_grad, nullptr); }

int exa_sample_points_stats(exa_renderer_t *r, const float *d_xyz, size_t n, int channel, ExaStats *stats)
{
  if (!stats) { setError("null stats"); return EXA_ERR_INVALID; }
  return samplePointsDev(r, d_xyz, n, channel, nullptr, nullptr, nullptr, stats);
}

// Host buffers in, host buffers out.  The caller's arrays are pageable, so the query is cut into chunks that travel
// through two sets of pinned staging buffers owned by the renderer: while chunk c is copied up, sampled and copied
// down by the GPU, the host copies chunk c-1 out of the other staging set and chunk c+1 into it (all cores).
static void hostCopyParallel(void *dst, const void *src, size_t n)
{
  if (n < (1u << 19)) { memcpy(dst, src, n); return; }
  const int T = std::max(1, std::min(omp_get_max_threads(), 16));
#pragma omp parallel for num_threads(T) schedule(static)
  for (int t = 0; t < T; t++) { const size_t b0 = n * t / T, e0 = n * (t + 1) / T; memcpy((char *)dst + b0, (const char *)src + b0, e0 - b0); }
}

int exa_sample_points_host(exa_renderer_t *r, const float *xyz, size_t n, int channel, float *values, uint8_t *valid, float *grad)
{
  if (!r) { setError("null renderer"); return EXA_ERR_INVALID; }
  if (n == 0) return EXA_OK;
  if (!xyz || !values || !valid) { setError("sample_points: null argument"); return EXA_ERR_INVALID; }
  if (channel < 0 || channel >= r->numFields) { setError("sample_points: channel out of range"); return EXA_ERR_INVALID; }
  CKR(cudaSetDevice(r->device));
  const size_t CH = exa_renderer::STAGE_POINTS;
  if (!r->h_stage[0]) {
    // per set: xyz (12 B) + value (4) + gradient (12) + valid (1) per point, device and pinned host
    for (int b = 0; b < 2; b++) {
      CKR(cudaMallocHost(&r->h_stage[b], CH * 29)); CKR(cudaMalloc(&r->d_stage[b], CH * 29));
      CKR(cudaEventCreateWithFlags(&r->evStage[b], cudaEventDisableTiming));
    }
  }
  auto hx = [&](int b) { return (float *)r->h_stage[b]; };        auto dx = [&](int b) { return (float *)r->d_stage[b]; };
  auto hv = [&](int b) { return (float *)(r->h_stage[b] + CH * 12); }; auto dv = [&](int b) { return (float *)(r->d_stage[b] + CH * 12); };
  auto hg = [&](int b) { return (float *)(r->h_stage[b] + CH * 16); }; auto dg = [&](int b) { return (float *)(r->d_stage[b] + CH * 16); };
  auto ho = [&](int b) { return (uint8_t *)(r->h_stage[b] + CH * 28); }; auto dok = [&](int b) { return (uint8_t *)(r->d_stage[b] + CH * 28); };
  auto drain = [&](size_t off, size_t cnt, int b) -> int {          // chunk finished on the device: pinned -> caller's arrays
    CKR(cudaEventSynchronize(r->evStage[b]));
    hostCopyParallel(values + off, hv(b), cnt * 4); hostCopyParallel(valid + off, ho(b), cnt);
    if (grad) hostCopyParallel(grad + 3 * off, hg(b), cnt * 12);
    return EXA_OK;
  };
  // chunks of the staging size: smaller ones lose more to the per-chunk calls than they gain in overlap (1 M points: 1.06 ms
  // with 262 144 points per chunk, 1.18 / 1.42 ms with 131 072 / 65 536)
  size_t step = CH;
  if (const char *e = getenv("EXA_POINTS_CHUNK")) step = std::min(CH, std::max<size_t>(1024, (size_t)atol(e)));   // experiment knob
  const size_t nChunks = (n + step - 1) / step;
  auto cntOf = [&](size_t c) { return std::min(step, n - c * step); };
  int rc = EXA_OK;
  hostCopyParallel(hx(0), xyz, cntOf(0) * 12);
  for (size_t c = 0; c < nChunks && !rc; c++) {
    const size_t off = c * step, cnt = cntOf(c); const int b = (int)(c & 1);
    CKR(cudaMemcpyAsync(dx(b), hx(b), cnt * 12, cudaMemcpyHostToDevice, r->stream));
    rc = samplePointsDev(r, dx(b), cnt, channel, dv(b), dok(b), grad ? dg(b) : nullptr, nullptr);
    if (rc) break;
    CKR(cudaMemcpyAsync(hv(b), dv(b), cnt * 4, cudaMemcpyDeviceToHost, r->stream));
    CKR(cudaMemcpyAsync(ho(b), dok(b), cnt, cudaMemcpyDeviceToHost, r->stream));
    if (grad) CKR(cudaMemcpyAsync(hg(b), dg(b), cnt * 12, cudaMemcpyDeviceToHost, r->stream));
    CKR(cudaEventRecord(r->evStage[b], r->stream));
    // while the GPU works on chunk c: results of chunk c-1 out of the other set (its event also says that the set's upload
    // is done), then the points of chunk c+1 into it
    if (c >= 1) rc = drain(off - step, cntOf(c - 1), b ^ 1);
    if (!rc && c + 1 < nChunks) hostCopyParallel(hx(b ^ 1), xyz + 3 * (off + step), cntOf(c + 1) * 12);
  }
  if (!rc) rc = drain((nChunks - 1) * step, cntOf(nChunks - 1), (int)((nChunks - 1) & 1));
  if (rc) cudaStreamSynchronize(r->stream);
  return rc;
}

// ---- tracer (OptixRenderer.cpp:444-487) ----------------------------------------------
int exa_renderer_set_tracer_enabled(exa_renderer_t *r, int enable) { if (!r) { setError("null renderer"); return EXA_ERR_INVALID; } r->tracerEnabled = enable != 0; return EXA_OK; }

int exa_renderer_set_tracer_params(exa_renderer_t *r, int numTraces, int numTimesteps, float steplen, const int channels[3],
                                   const float seedLo[3], const float seedHi[3])
{
  if (!r || numTraces < 0 || numTimesteps < 1) { setError("set_tracer_params: bad argument"); return EXA_ERR_INVALID; }
  if (channels) for (int i = 0; i < 3; i++) if (channels[i] < 0 || channels[i] >= r->numFields) { setError("tracer channel out of range"); return EXA_ERR_INVALID; }
  // the shape of d_traces changes only where the buffer is rebuilt (reset_tracer / set_tracer_seeds / set_traces), as in
  // the reference, which pushes these values to the device inside resetTracer (OptixRenderer.cpp:450-474)
  r->pendNumTraces = numTraces; r->pendNumTimesteps = numTimesteps; r->steplen = steplen;
  if (channels) memcpy(r->tracerChannels, channels, 12);
  if (seedLo) memcpy(r->seedLo, seedLo, 12);
  if (seedHi) memcpy(r->seedHi, seedHi, 12);
  return EXA_OK;
}

static int uploadSeeds(exa_renderer *r, const std::vector<float> &host)
{
  CKR(cudaSetDevice(r->device));
  const size_t n = host.size();
  if (n > r->tracesCap) { CKR(cudaStreamSynchronize(r->stream)); cudaFree(r->d_traces); r->d_traces = nullptr; CKR(cudaMalloc(&r->d_traces, std::max(n, (size_t)3) * 4)); r->tracesCap = n; }
  if (n) CKR(cudaMemcpyAsync(r->d_traces, host.data(), n * 4, cudaMemcpyHostToDevice, r->stream));
  CKR(cudaStreamSynchronize(r->stream));
  r->timestep = 0; r->traceBoundsDirty = true;
  return EXA_OK;
}

int exa_renderer_reset_tracer(exa_renderer_t *r)
{
  if (!r) { setError("null renderer"); return EXA_ERR_INVALID; }
  for (int i = 0; i < 3; i++) if (r->tracerChannels[i] >= r->numFields) { setError("tracer needs 3 scalar channels (vx,vy,vz)"); return EXA_ERR_INVALID; }
  // OptixRenderer::resetTracer (:450-460): seeds from std::default_random_engine(0), x,y,z in
  // that order, uniform in seedRegion * voxel-bounds EXTENT (not offset by bounds.lower)
  r->numTraces = r->pendNumTraces; r->numTimesteps = r->pendNumTimesteps;
  std::vector<float> host((size_t)r->numTraces * r->numTimesteps * 3, 0.f);
  const float size[3] = {r->voxelHi[0] - r->voxelLo[0], r->voxelHi[1] - r->voxelLo[1], r->voxelHi[2] - r->voxelLo[2]};
  std::default_random_engine engine(0);
  std::uniform_real_distribution<float> x(r->seedLo[0] * size[0], r->seedHi[0] * size[0]);
  std::uniform_real_distribution<float> y(r->seedLo[1] * size[1], r->seedHi[1] * size[1]);
  std::uniform_real_distribution<float> z(r->seedLo[2] * size[2], r->seedHi[2] * size[2]);
  for (int i = 0; i < r->numTraces; i++) {
    float *p = host.data() + (size_t)i * r->numTimesteps * 3;
    p[0] = x(engine); p[1] = y(engine); p[2] = z(engine);
  }
  return uploadSeeds(r, host);
}

int exa_renderer_set_tracer_seeds(exa_renderer_t *r, const float *xyz, int numTraces)
{
  if (!r || !xyz || numTraces < 0) { setError("set_tracer_seeds: bad argument"); return EXA_ERR_INVALID; }
  for (int i = 0; i < 3; i++) if (r->tracerChannels[i] >= r->numFields) { setError("tracer needs 3 scalar channels (vx,vy,vz)"); return EXA_ERR_INVALID; }
  r->numTraces = r->pendNumTraces = numTraces; r->numTimesteps = r->pendNumTimesteps;
  std::vector<float> host((size_t)numTraces * r->numTimesteps * 3, 0.f);
  for (int i = 0; i < numTraces; i++) memcpy(host.data() + (size_t)i * r->numTimesteps * 3, xyz + 3 * i, 12);
  return uploadSeeds(r, host);
}

static int traceSteps(exa_renderer *r, int first, int last)
{
  if (!r->d_traces) { setError("tracer: resetTracer was never called"); return EXA_ERR_INVALID; }
  CKR(cudaSetDevice(r->device));
  RenderParams RP; fillParams(r, RP);
  TraceParams P; P.scene = RP.scene; P.traces = r->d_traces; P.numTraces = r->numTraces; P.numTimesteps = r->numTimesteps;
  P.firstStep = first; P.lastStep = last; memcpy(P.channels, r->tracerChannels, 12); P.steplen = r->steplen;
  memcpy(P.worldLo, r->worldLo, 12); memcpy(P.worldHi, r->worldHi, 12);
  const int k = launchTrace(P, r->stream);
  if (k < 0) { setError(std::string("trace kernel launch failed: ") + cudaGetErrorString(cudaGetLastError())); return EXA_ERR_CUDA; }
  r->launches += k;
  return EXA_OK;
}

int exa_renderer_advance_tracer(exa_renderer_t *r, int *rebuilt)
{
  if (!r) { setError("null renderer"); return EXA_ERR_INVALID; }
  if (rebuilt) *rebuilt = 0;
  if (!r->tracerEnabled) return EXA_OK;                      // advanceTracer (:476-487)
  r->timestep++; r->traceBoundsDirty = true;
  if (r->timestep <= r->numTimesteps) {
    if (rebuilt) *rebuilt = 1;
    if (r->timestep < r->numTimesteps) return traceSteps(r, r->timestep, r->timestep);   // device: t < numTimesteps (exabrick.cu:1539)
  }
  return EXA_OK;
}

int exa_renderer_trace_all(exa_renderer_t *r)
{
  if (!r) { setError("null renderer"); return EXA_ERR_INVALID; }
  const int first = r->timestep + 1;
  if (first >= r->numTimesteps) return EXA_OK;
  int rc = traceSteps(r, first, r->numTimesteps - 1);
  if (!rc) r->timestep = r->numTimesteps;
  r->traceBoundsDirty = true;
  return rc;
}

int exa_renderer_set_traces(exa_renderer_t *r, const float *xyz, int numTraces, int numTimesteps, int currentTimestep)
{
  if (!r || !xyz || numTraces < 0 || numTimesteps < 1 || currentTimestep < 0) { setError("set_traces: bad argument"); return EXA_ERR_INVALID; }
  r->numTraces = r->pendNumTraces = numTraces; r->numTimesteps = r->pendNumTimesteps = numTimesteps;
  std::vector<float> host(xyz, xyz + (size_t)numTraces * numTimesteps * 3);
  int rc = uploadSeeds(r, host);
  if (rc) return rc;
  r->timestep = currentTimestep; r->traceBoundsDirty = true;
  return EXA_OK;
}

int exa_renderer_read_traces(exa_renderer_t *r, float *xyz)
{
  if (!r || !xyz || !r->d_traces) { setError("read_traces: no traces"); return EXA_ERR_INVALID; }
  CKR(cudaSetDevice(r->device));
  CKR(cudaMemcpyAsync(xyz, r->d_traces, (size_t)r->numTraces * r->numTimesteps * 12, cudaMemcpyDeviceToHost, r->stream));
  CKR(cudaStreamSynchronize(r->stream));
  return EXA_OK;
}

// ---- multi-GPU ------------------------------------------------------------------------
int exa_renderer_set_partition(exa_renderer_t *r, int mode, int rank, int numRanks)
{
  if (!r || numRanks < 1 || rank < 0 || rank >= numRanks) { setError("set_partition: bad rank"); return EXA_ERR_INVALID; }
  if (mode != EXA_PARTITION_NONE && mode != EXA_PARTITION_SORT_LAST && mode != EXA_PARTITION_IMAGE_TILES) { setError("set_partition: bad mode"); return EXA_ERR_INVALID; }
  if (numRanks > EXA_MAX_PARTS && mode == EXA_PARTITION_SORT_LAST) { setError("set_partition: at most 16 sort-last parts"); return EXA_ERR_LIMIT; }
  if (r->dataPartitioned) {
    // the scene of this renderer IS one part (exa_renderer_create_partitioned): nothing else is resident
    if (mode == EXA_PARTITION_SORT_LAST && rank == r->rank && numRanks == r->numRanks) return EXA_OK;
    setError("set_partition: this renderer holds only the bricks of its own sort-last part"); return EXA_ERR_INVALID;
  }
  r->partMode = mode; r->rank = rank; r->numRanks = numRanks; r->masksDirty = true;
  r->tileOrderValid = false; r->tileListLen = 0;             // the set of owned tiles may have changed
  r->partTree.clear(); r->partRoot = 0;
  if (mode != EXA_PARTITION_SORT_LAST) return EXA_OK;
  // replicated scene: the same partition tree as the data-partitioned mode, applied as a visibility mask
  CKR(cudaSetDevice(r->device));
  std::vector<ExaBrickRegion> regions(r->abr.numRegions); std::vector<int32_t> leafList(r->abr.leafListLen);
  std::vector<ExaBrick> hb(r->numBricks);
  CKR(cudaStreamSynchronize(r->stream));
  if (!regions.empty()) CKR(cudaMemcpy(regions.data(), r->abr.regions, sizeof(ExaBrickRegion) * regions.size(), cudaMemcpyDeviceToHost));
  if (!leafList.empty()) CKR(cudaMemcpy(leafList.data(), r->abr.leafList, 4 * leafList.size(), cudaMemcpyDeviceToHost));
  CKR(cudaMemcpy(hb.data(), r->d_bricks, sizeof(ExaBrick) * hb.size(), cudaMemcpyDeviceToHost));
  std::vector<double> cost;
  const char *bal = getenv("EXA_PARTITION_BALANCE");
  nodeCosts(r->h_kd, regions, leafList, hb.data(), bal && !strcmp(bal, "cells"), cost);
  buildPartitionTree(r, numRanks, cost);
  r->partRoot = partitionRootOf(r, rank);
  const int k = launchPartitionMask(r->abr.kd, r->abr.levelOffset, r->partRoot, r->d_inPart, r->stream);
  if (k < 0) { setError("partition mask kernels failed"); return EXA_ERR_CUDA; }
  r->launches += k;
  return EXA_OK;
}

int exa_renderer_partition_order(exa_renderer_t *r, int *order, int *numParts)
{
  if (!r || !order || r->partMode != EXA_PARTITION_SORT_LAST || r->partTree.empty()) { setError("partition_order: renderer is not in sort-last mode"); return EXA_ERR_INVALID; }
  // camera position in voxel space decides, at every plane of the partition tree, which side comes first
  // for EVERY pixel (all primary rays start there); ranks beyond the number of parts own nothing and go last
  const float *m = (const float *)&r->fs.voxelSpaceTransform; const float *p = r->fs.camera.pos;
  float cam[3];
  for (int a = 0; a < 3; a++) cam[a] = p[0] * m[a] + p[1] * m[3 + a] + p[2] * m[6 + a] + m[9 + a];
  int n = 0;
  std::vector<int> st; st.push_back(0);
  while (!st.empty()) {
    const int i = st.back(); st.pop_back();
    const exa_renderer::PartNode &pn = r->partTree[i];
    if (pn.left < 0) { order[n++] = pn.rank; continue; }
    const KdNode nd = r->h_kd[pn.kd];
    const int axis = (int)(nd.info >> 30);
    const bool camBelow = cam[axis] < nd.split;              // left child = below the plane
    const int first = camBelow ? pn.left : pn.right, second = camBelow ? pn.right : pn.left;
    st.push_back(second); st.push_back(first);
  }
  if (numParts) *numParts = n;
  for (int k = n; k < r->numRanks; k++) order[k] = k;        // empty parts
  return EXA_OK;
}

int exa_renderer_render_partial(exa_renderer_t *r, float *d_partialRGBA, float *d_partialSurf)
{
  if (!d_partialRGBA) { setError("null partial buffer"); return EXA_ERR_INVALID; }
  return doRender(r, (float4 *)d_partialRGBA, (float4 *)d_partialSurf, nullptr, false);
}

int exa_renderer_composite_partials(exa_renderer_t *r, const float *const *d_partials, const float *const *d_surfs, int numParts,
                                    size_t first, size_t count, uint32_t *d_rgba8Out)
{
  if (!r || !d_partials || !d_rgba8Out || !r->d_accum) { setError("composite_partials: bad argument"); return EXA_ERR_INVALID; }
  if (numParts < 1 || numParts > EXA_MAX_PARTS) { setError("composite_partials: 1..16 parts"); return EXA_ERR_INVALID; }
  if (first + count > (size_t)r->width * r->height) { setError("composite_partials: span outside the frame"); return EXA_ERR_INVALID; }
  CKR(cudaSetDevice(r->device));
  RenderParams P; fillParams(r, P);
  CompositeParams C; memset(&C, 0, sizeof(C));
  for (int j = 0; j < numParts; j++) {
    if (!d_partials[j] || (d_surfs && !d_surfs[j])) { setError("composite_partials: null partial"); return EXA_ERR_INVALID; }
    C.part[j] = (const float4 *)d_partials[j]; C.surf[j] = d_surfs ? (const float4 *)d_surfs[j] : nullptr;
  }
  C.n = numParts; C.hasSurf = d_surfs != nullptr; C.first = first; C.count = count; C.out = d_rgba8Out;
  // what every part wrote: the rectangle of its box (doRender), the whole frame when replicated surfaces are on or the
  // renderer is not in sort-last mode (buffers filled by the caller)
  for (int j = 0; j < numParts; j++) { C.rect[j][0] = 0; C.rect[j][1] = 0; C.rect[j][2] = r->width - 1; C.rect[j][3] = r->height - 1; }
  if (sortLastParts(r) && P.anySurf != 2 && r->tileListLen == 0) {
    int order[EXA_MAX_PARTS], np = 0;
    const int rc = exa_renderer_partition_order(r, order, &np); if (rc) return rc;
    for (int j = 0; j < numParts && j < r->numRanks; j++) {
      float lo[3], hi[3];
      if (partBoxOf(r, order[j], lo, hi)) boxScreenRect(r, lo, hi, C.rect[j]);
      else { C.rect[j][0] = C.rect[j][1] = 0; C.rect[j][2] = C.rect[j][3] = -1; }
    }
  }
  const int k = launchCompositeFinish(P, C, r->stream);
  if (k < 0) { setError("composite kernel launch failed"); return EXA_ERR_CUDA; }
  r->launches += k; return EXA_OK;
}

// ---- image tiles balanced by measured cost --------------------------------------------------------------
int exa_renderer_enable_tile_costs(exa_renderer_t *r, int enable)
{
  if (!r || !r->d_accum) { setError("enable_tile_costs: resizeFrameBuffer was never called"); return EXA_ERR_INVALID; }
  CKR(cudaSetDevice(r->device));
  const int numTiles = ((r->width + 15) / 16) * ((r->height + 7) / 8);
  if (!enable) { CKR(cudaStreamSynchronize(r->stream)); cudaFree(r->d_tileCost); r->d_tileCost = nullptr; r->tileCostLen = 0; return EXA_OK; }
  if (!r->d_tileCost) { CKR(cudaMalloc(&r->d_tileCost, 4 * (size_t)numTiles)); r->tileCostLen = numTiles; }
  CKR(cudaMemsetAsync(r->d_tileCost, 0, 4 * (size_t)numTiles, r->stream));
  return EXA_OK;
}

int exa_renderer_get_tile_costs(exa_renderer_t *r, uint32_t *costs, int numTiles)
{
  if (!r || !costs || !r->d_tileCost || numTiles != r->tileCostLen) { setError("get_tile_costs: tile costs are not enabled / wrong tile count"); return EXA_ERR_INVALID; }
  CKR(cudaSetDevice(r->device));
  CKR(cudaMemcpyAsync(costs, r->d_tileCost, 4 * (size_t)numTiles, cudaMemcpyDeviceToHost, r->stream));
  CKR(cudaStreamSynchronize(r->stream));
  return EXA_OK;
}

int exa_renderer_set_tile_list(exa_renderer_t *r, const uint32_t *tiles, int numTiles)
{
  if (!r || !r->d_accum || numTiles < 0 || (numTiles && !tiles)) { setError("set_tile_list: bad argument"); return EXA_ERR_INVALID; }
  const int total = ((r->width + 15) / 16) * ((r->height + 7) / 8);
  for (int i = 0; i < numTiles; i++) if (tiles[i] >= (uint32_t)total) { setError("set_tile_list: tile id outside the frame"); return EXA_ERR_INVALID; }
  CKR(cudaSetDevice(r->device));
  if ((size_t)numTiles > r->tileListCap) {
    CKR(cudaStreamSynchronize(r->stream)); cudaFree(r->d_tileList); r->d_tileList = nullptr;
    CKR(cudaMalloc(&r->d_tileList, 4 * (size_t)numTiles)); r->tileListCap = numTiles;
  }
  if (numTiles) { CKR(cudaMemcpyAsync(r->d_tileList, tiles, 4 * (size_t)numTiles, cudaMemcpyHostToDevice, r->stream)); CKR(cudaStreamSynchronize(r->stream)); }
  r->tileListLen = numTiles;
  return EXA_OK;
}

int exa_balance_tiles(const uint32_t *costs, int numTiles, int numRanks, int rank, uint32_t *outTiles, int *outCount)
{
  if (!costs || !outTiles || !outCount || numTiles < 0 || numRanks < 1 || rank < 0 || rank >= numRanks) { setError("balance_tiles: bad argument"); return EXA_ERR_INVALID; }
  // longest-processing-time-first: tiles by descending cost (ties: ascending id), each to the rank with the smallest load
  // (ties: the lowest rank).  Deterministic, so every rank computes the same deal from the same costs.
  std::vector<uint32_t> order(numTiles);
  for (int i = 0; i < numTiles; i++) order[i] = (uint32_t)i;
  std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return costs[a] > costs[b]; });
  std::vector<unsigned long long> load(numRanks, 0); std::vector<int> cnt(numRanks, 0);
  int n = 0;
  const int cap = (numTiles + numRanks - 1) / numRanks + numTiles / (4 * numRanks) + 1;   // no rank gets more than ~1.25x its share of tiles (zero-cost tails)
  for (int i = 0; i < numTiles; i++) {
    int best = -1;
    for (int k = 0; k < numRanks; k++) if (cnt[k] < cap && (best < 0 || load[k] < load[best])) best = k;
    load[best] += (unsigned long long)costs[order[i]] + 1ull; cnt[best]++;
    if (best == rank) outTiles[n++] = order[i];
  }
  *outCount = n;
  return EXA_OK;
}

int exa_peer_barrier(void *const *flagPtrs, int rank, int numRanks, unsigned seq, void *cudaStream)
{
  if (!flagPtrs || numRanks < 1 || numRanks > EXA_MAX_PARTS || rank < 0 || rank >= numRanks) { setError("peer_barrier: bad argument"); return EXA_ERR_INVALID; }
  PeerBarrierParams B; memset(&B, 0, sizeof(B));
  for (int k = 0; k < numRanks; k++) { if (!flagPtrs[k]) { setError("peer_barrier: null flag array"); return EXA_ERR_INVALID; } B.flags[k] = (unsigned *)flagPtrs[k]; }
  B.rank = rank; B.n = numRanks; B.seq = seq; B.timeoutCycles = 4000000000ll;      // ~2 s at 1.9 GHz
  B.err = (unsigned *)flagPtrs[rank] + EXA_MAX_PARTS;                              // word 16 of the local array: sticky time-out flag
  if (launchPeerBarrier(B, (cudaStream_t)cudaStream) < 0) { setError(std::string("peer barrier launch failed: ") + cudaGetErrorString(cudaGetLastError())); return EXA_ERR_CUDA; }
  return EXA_OK;
}

int exa_copy_async(void *dst, const void *src, size_t bytes, void *cudaStream)
{
  if (!dst || !src) { setError("copy_async: null pointer"); return EXA_ERR_INVALID; }
  CKR(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, (cudaStream_t)cudaStream));
  return EXA_OK;
}

int exa_host_register(void *ptr, size_t bytes)
{
  if (!ptr || !bytes) { setError("host_register: bad argument"); return EXA_ERR_INVALID; }
  CKR(cudaHostRegister(ptr, bytes, cudaHostRegisterPortable));
  return EXA_OK;
}
int exa_host_unregister(void *ptr) { if (ptr) CKR(cudaHostUnregister(ptr)); return EXA_OK; }

} // extern "C"
