// 2-D decomposition (params.nslabs_y > 1): ranges of rows times z-slabs.  Included by heat3d.cu inside its anonymous
// namespace, after the single-axis machinery it builds on.
//
// The reference cuts the domain along all three axes (MPI_Dims_create / MPI_Cart_create, heat3D.cu:243-263, chunk
// sizes :385-389) and packs the faces that are strided in its memory layout element by element on the host
// (:562-605).  Here the x axis stays whole (it is the contiguous one: cutting it would break the 128-byte rows every
// kernel streams), z is cut into slabs as everywhere else in this library, and y is cut into `nslabs_y` ranges of rows.
// A y-face is one row of every plane - nx contiguous doubles, then a stride of a whole plane - and moves through the
// K3 kernels of kernels.cuh: pack into a dense (plane, x) buffer, one copy to the neighbour's device, unpack into its
// ghost row.
//
// Structure: the context the caller holds is a PARENT that owns one ordinary child context per row range.  A child is
// told it is a domain of `rows + ghost rows` rows; its first and last local rows are cells its kernels never write -
// the true Dirichlet rows at the ends of the domain, one ghost row where a neighbouring range begins - and the parent
// refreshes the ghost rows from the neighbours' outermost owned rows after every update.  Everything else (z-slabs,
// their overlapped exchange, residual, error norm, init) is the child's ordinary code, evaluated at global row
// coordinates through the child's y offset.  One update per exchange: temporal_k is 1 here (depth 2 would need the
// corner cells that are ghost in y and z at once).  Ownership is non-overlapping, so the field is the single-domain
// field bit for bit; the reference's shared interface planes with extrapolated edges and averaged corners
// (:861-1011) are not reproduced.

int g2_fail_unsupported(const char* what)
{
    return fail(HEAT3D_ERR_INVALID, "%s is not available on a context with nslabs_y > 1", what);
}

void g2_destroy(heat3d_ctx* c)
{
    for (heat3d_ctx::YLink& L : c->ylinks) {
        if (L.from_kid < (int)c->kids.size() && c->kids[L.from_kid]) cudaSetDevice(c->kids[L.from_kid]->slabs[L.slab].dev);
        if (L.stage_src) cudaFree(L.stage_src);
        if (L.ev_packed) cudaEventDestroy(L.ev_packed);
        if (L.stage_dst && L.stage_dst != L.stage_src) {
            cudaSetDevice(c->kids[L.to_kid]->slabs[L.slab].dev);
            cudaFree(L.stage_dst);
        }
        if (L.ev_unpacked) cudaEventDestroy(L.ev_unpacked);
    }
    c->ylinks.clear();
    if (!c->kids.empty() && c->kids[0]) cudaSetDevice(c->kids[0]->slabs[0].dev);
    if (c->g2_t0) cudaEventDestroy(c->g2_t0);
    if (c->g2_t1) cudaEventDestroy(c->g2_t1);
    c->g2_t0 = c->g2_t1 = nullptr;
    for (heat3d_ctx* k : c->kids) heat3d_destroy(k);
    c->kids.clear();
}

int g2_create(const heat3d_params& p, heat3d_ctx** out)
{
    const int py = p.nslabs_y, pz = p.nslabs > 0 ? p.nslabs : 1;
    if (p.temporal_k > 1) return fail(HEAT3D_ERR_INVALID, "nslabs_y > 1 takes temporal_k = 1 (got %d)", p.temporal_k);
    if (p.nlocal != 0 && p.nlocal != pz) return fail(HEAT3D_ERR_INVALID, "nslabs_y > 1 is single-process: nlocal must be 0 or nslabs");
    if (p.slab_first != 0 || p.nccl_unique_id) return fail(HEAT3D_ERR_INVALID, "nslabs_y > 1 is single-process");
    if ((uint32_t)py > p.ny / 2) return fail(HEAT3D_ERR_INVALID, "%d row ranges for %u rows: every range needs >= 2 rows", py, p.ny);
    if (p.stream) return fail(HEAT3D_ERR_INVALID, "a caller-provided stream needs a single slab");
    heat3d_ctx* c = new (std::nothrow) heat3d_ctx();
    if (!c) return fail(HEAT3D_ERR_ALLOC, "out of host memory");
    c->p = p;
    c->p.devices = nullptr;
    memset(&c->tm, 0, sizeof c->tm);
    int cur_dev = 0;
    cudaGetDevice(&cur_dev);
    std::vector<int32_t> devs((size_t)py * pz);
    for (size_t i = 0; i < devs.size(); ++i) devs[i] = p.devices ? p.devices[i] : cur_dev;
    for (int iy = 0; iy < py; ++iy) {
        const uint32_t y0 = (uint32_t)((uint64_t)iy * p.ny / py), y1 = (uint32_t)((uint64_t)(iy + 1) * p.ny / py);
        const uint32_t glo = iy > 0 ? 1u : 0u, ghi = iy < py - 1 ? 1u : 0u;
        heat3d_params q = p;
        q.nslabs_y = 0;
        q.ny = (y1 - y0) + glo + ghi;
        q.nslabs = pz; q.slab_first = 0; q.nlocal = pz;
        q.devices = &devs[(size_t)iy * pz];
        heat3d_ctx* k = nullptr;
        const int rc = heat3d_create(&q, &k);
        if (rc) { g2_destroy(c); delete c; return rc; }
        k->y_off = (int)(y0 - glo);
        k->ny_global = p.ny;
        k->want_raw_error_sum = true;
        c->kids.push_back(k);
    }
    // faces between neighbouring row ranges: per z-slab one link in each direction
    for (int iy = 0; iy + 1 < py; ++iy)
        for (int i = 0; i < pz; ++i)
            for (int dir = 0; dir < 2; ++dir) {
                heat3d_ctx::YLink L;
                L.slab = i;
                if (dir == 0) { L.from_kid = iy; L.to_kid = iy + 1; L.src_row = (int)c->kids[iy]->p.ny - 2; L.dst_row = 0; }
                else { L.from_kid = iy + 1; L.to_kid = iy; L.src_row = 1; L.dst_row = (int)c->kids[iy]->p.ny - 1; }
                Slab& sf = c->kids[L.from_kid]->slabs[i];
                Slab& st = c->kids[L.to_kid]->slabs[i];
                const size_t bytes = (size_t)(sf.z1 - sf.z0) * p.nx * sizeof(double);
                cudaError_t e = cudaSetDevice(sf.dev);
                if (e == cudaSuccess) e = cudaMalloc(&L.stage_src, bytes);
                if (e == cudaSuccess) e = cudaEventCreateWithFlags(&L.ev_packed, cudaEventDisableTiming);
                L.stage_dst = L.stage_src;
                if (e == cudaSuccess && st.dev != sf.dev) {
                    e = cudaSetDevice(st.dev);
                    if (e == cudaSuccess) e = cudaMalloc(&L.stage_dst, bytes);
                }
                if (e == cudaSuccess) e = cudaSetDevice(st.dev);
                if (e == cudaSuccess) e = cudaEventCreateWithFlags(&L.ev_unpacked, cudaEventDisableTiming);
                c->ylinks.push_back(L);
                if (e != cudaSuccess) {
                    fail(HEAT3D_ERR_ALLOC, "setting up the face buffers of the y split failed: %s", cudaGetErrorString(e));
                    g2_destroy(c); delete c;
                    return HEAT3D_ERR_ALLOC;
                }
            }
    cudaSetDevice(c->kids[0]->slabs[0].dev);
    if (cudaEventCreate(&c->g2_t0) != cudaSuccess || cudaEventCreate(&c->g2_t1) != cudaSuccess) {
        fail(HEAT3D_ERR_CUDA, "cudaEventCreate failed");
        g2_destroy(c); delete c;
        return HEAT3D_ERR_CUDA;
    }
    c->g2_exchange = std::string("y: pack/unpack kernels + copy; z: ") + heat3d_exchange_name(c->kids[0]);
    *out = c;
    return HEAT3D_OK;
}

// Refresh every ghost row of the children's CURRENT buffers from the neighbouring ranges' outermost owned rows.
// pack on the sender's compute stream (after its stencil launch, by stream order) -> copy -> unpack on the receiver's
// compute stream (before its next stencil launch, by stream order).
int g2_exchange_rows(heat3d_ctx* c)
{
    const int nx = (int)c->p.nx;
    for (heat3d_ctx::YLink& L : c->ylinks) {
        heat3d_ctx* kf = c->kids[L.from_kid];
        heat3d_ctx* kt = c->kids[L.to_kid];
        Slab& sf = kf->slabs[L.slab];
        Slab& st = kt->slabs[L.slab];
        const int nzl = (int)(sf.z1 - sf.z0);
        const size_t bytes = (size_t)nzl * nx * sizeof(double);
        int rc = set_dev(sf);
        if (rc) return rc;
        CU_TRY(cudaStreamWaitEvent(sf.compute, L.ev_unpacked, 0));       // the staging buffer's previous load has been unpacked
        for (int p0 = 0; p0 < nzl; p0 += 65535) {
            const int n = std::min(65535, nzl - p0);
            h3::face_rows_kernel<true><<<dim3(1, n), 256, 0, sf.compute>>>(sf.buf[kf->cur], L.stage_src + (size_t)p0 * nx, nx, 1, L.src_row,
                                                                         (int)sf.G + p0, (int)sf.pitch, sf.plane);
            CU_TRY(cudaGetLastError());
            c->tm.launches++; c->tm.last_call_launches++;
        }
        if (L.stage_dst != L.stage_src) CU_TRY(cudaMemcpyPeerAsync(L.stage_dst, st.dev, L.stage_src, sf.dev, bytes, sf.compute));
        CU_TRY(cudaEventRecord(L.ev_packed, sf.compute));
        if ((rc = set_dev(st))) return rc;
        CU_TRY(cudaStreamWaitEvent(st.compute, L.ev_packed, 0));
        for (int p0 = 0; p0 < nzl; p0 += 65535) {
            const int n = std::min(65535, nzl - p0);
            h3::face_rows_kernel<false><<<dim3(1, n), 256, 0, st.compute>>>(st.buf[kt->cur], L.stage_dst + (size_t)p0 * nx, nx, 1, L.dst_row,
                                                                          (int)st.G + p0, (int)st.pitch, st.plane);
            CU_TRY(cudaGetLastError());
            c->tm.launches++; c->tm.last_call_launches++;
        }
        CU_TRY(cudaEventRecord(L.ev_unpacked, st.compute));
    }
    return HEAT3D_OK;
}

int g2_sync(heat3d_ctx* c)
{
    for (heat3d_ctx* k : c->kids) {
        const int rc = heat3d_sync(k);
        if (rc) return rc;
    }
    return HEAT3D_OK;
}

// nsteps updates; with `residual` the last one reports max |T_new - T_old| over all children
int g2_step(heat3d_ctx* c, uint32_t nsteps, double* residual)
{
    Slab& s0 = c->kids[0]->slabs[0];
    int rc = set_dev(s0);
    if (rc) return rc;
    c->tm.last_call_launches = 0;
    c->tm.last_call_stencil_launches = 0;
    CU_TRY(cudaEventRecord(c->g2_t0, s0.compute));           // call bracket on one stream: the ranges are coupled every update
    double worst = 0.0;
    for (uint32_t i = 0; i < nsteps; ++i) {
        const bool last = residual && i + 1 == nsteps;
        for (heat3d_ctx* k : c->kids) {
            if (last) {
                double r = 0.0;
                if ((rc = heat3d_step_residual(k, 1, &r))) return rc;
                worst = std::max(worst, r);
            } else if ((rc = heat3d_step(k, 1))) return rc;
            c->tm.launches += k->tm.last_call_launches;
            c->tm.last_call_launches += k->tm.last_call_launches;
            c->tm.last_call_stencil_launches += k->tm.last_call_stencil_launches;
        }
        if ((rc = g2_exchange_rows(c))) return rc;
        c->tm.steps++;
    }
    if ((rc = set_dev(s0))) return rc;
    CU_TRY(cudaEventRecord(c->g2_t1, s0.compute));
    c->timing_pending = true;
    if (residual) {
        *residual = worst;
        c->have_res = true;
        c->last_res = worst;
    } else if (nsteps)
        c->have_res = false;
    return HEAT3D_OK;
}

int g2_finish_timing(heat3d_ctx* c)
{
    if (!c->timing_pending) return HEAT3D_OK;
    int rc = g2_sync(c);
    if (rc) return rc;
    Slab& s0 = c->kids[0]->slabs[0];
    if ((rc = set_dev(s0))) return rc;
    float ms = 0.f;
    CU_TRY(cudaEventElapsedTime(&ms, c->g2_t0, c->g2_t1));
    c->tm.last_call_ms = ms;
    c->tm.total_ms += ms;
    c->timing_pending = false;
    return HEAT3D_OK;
}

int g2_init(heat3d_ctx* c, double sp_y)
{
    for (heat3d_ctx* k : c->kids) {
        const int rc = heat3d_init_reference_field(k, sp_y);      // evaluated at global rows through the child's y offset
        if (rc) return rc;
    }
    c->have_res = false;
    return HEAT3D_OK;
}

// Host <-> children for global planes [z0, z0 + n): each child holds rows [y_off, y_off + its ny) of every plane.
// The rows are gathered / scattered through a temporary host array in the child's own dense layout.
int g2_transfer(heat3d_ctx* c, double* host, uint32_t z0, uint32_t n, bool upload)
{
    const size_t nx = c->p.nx, ny = c->p.ny;
    for (size_t iy = 0; iy < c->kids.size(); ++iy) {
        heat3d_ctx* k = c->kids[iy];
        const size_t nyl = k->p.ny, yo = (size_t)k->y_off;
        // on the way back only the rows the child owns are taken (its ghost rows are copies of the neighbours' rows)
        const size_t own_lo = iy > 0 ? 1 : 0, own_hi = iy + 1 < c->kids.size() ? nyl - 1 : nyl;
        std::vector<double> tmp((size_t)n * nyl * nx);
        int rc;
        if (upload) {
            for (size_t z = 0; z < n; ++z)
                memcpy(&tmp[z * nyl * nx], host + (z * ny + yo) * nx, nyl * nx * sizeof(double));
            if ((rc = heat3d_upload_planes(k, tmp.data(), z0, n))) return rc;
        } else {
            if ((rc = heat3d_download_planes(k, tmp.data(), z0, n))) return rc;
            for (size_t z = 0; z < n; ++z)
                memcpy(host + (z * ny + yo + own_lo) * nx, &tmp[(z * nyl + own_lo) * nx], (own_hi - own_lo) * nx * sizeof(double));
        }
    }
    if (upload) c->have_res = false;
    return HEAT3D_OK;
}

int g2_error(heat3d_ctx* c, double sp_y, double* err)
{
    double sum = 0.0;
    for (heat3d_ctx* k : c->kids) {
        double part = 0.0;
        const int rc = heat3d_mean_abs_error_vs_y(k, sp_y, &part);     // children return their plain sums
        if (rc) return rc;
        sum += part;
    }
    const unsigned denom = (c->p.nx - 2u) * (c->p.ny - 2u) * (c->p.nz - 2u);      // the reference's `unsigned` divisor, cu:1102
    *err = sum / denom;
    return HEAT3D_OK;
}
