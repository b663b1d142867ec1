// pbe_mesh.h -- the device-side mesh handle shared by pbe_transport.cu (explicit upwind update) and
// pbe_mules.cu (MULES-limited update): OpenFOAM face addressing turned into cell-centred tables.
#pragma once

#include <vector>

#include "pbe_common.h"

struct pbe_mesh {
    pbe_ctx *ctx = nullptr;
    int64_t ncells = 0, nint = 0, nbnd = 0, nirr = 0;
    int maxf = 0; // most faces on any cell (info)
    int F = 0;    // face count of the regular cells (0: none)
    int32_t *d_other = nullptr; // [F][ncells]: neighbour cell, or -(b + 2) for boundary face b; row 0 holds -1 for irregular cells
    int32_t *d_face = nullptr;  // [F][ncells]: face index f (this cell owns it), ~f (it is the neighbour), nint + b (boundary)
    // host copies of the face addressing: pbe_mesh_set_coupled re-splits the cells
    std::vector<int32_t> h_owner, h_neigh, h_bown;
    double *d_vcell = nullptr;  // [ncells]
    // irregular cells, CSR: entries of cell irr_cell[i] are irr_ptr[i] .. irr_ptr[i+1]-1
    int32_t *d_irr_cell = nullptr, *d_irr_ptr = nullptr;
    int32_t *d_irr_other = nullptr; // >= 0 neighbour cell; -(b + 2) boundary face b
    int32_t *d_irr_face = nullptr;  // as d_face (boundary faces: n_internal + b)
    int64_t coupled0 = 0, ncoupled = 0; // boundary faces [coupled0, coupled0 + ncoupled): processor patches
    // the irregular cells' kernel runs beside the regular cells' one (disjoint outputs): fork / join
    cudaStream_t side = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    // resident fields of the host-buffer entry (pbe_moc_correct_host), grown on demand
    double *res = nullptr;
    size_t res_doubles = 0;
    // ---- MULES path (pbe_mules.cu) ------------------------------------------------------------------
    // CSR over all cells, ascending face order: entries of cell c are csr_ptr[c] .. csr_ptr[c+1]-1;
    // csr_other >= 0: the cell across an internal face, -(b + 2): boundary face b; csr_face: f (this
    // cell owns it), ~f (it is the neighbour), nint + b for boundary faces
    int32_t *d_csr_ptr = nullptr, *d_csr_other = nullptr, *d_csr_face = nullptr;
    int32_t *d_owner = nullptr, *d_neigh = nullptr, *d_bown = nullptr; // face -> cells
    // geometry for the limited schemes (pbe_mesh_set_geometry): Sf [3][nint + nbnd], C [3][ncells]
    // (component-major for coalescing), linear weights [nint]
    double *d_Sf = nullptr, *d_C = nullptr, *d_w = nullptr;
    double *mules_ws = nullptr; // workspace of pbe_transport_mules, grown on demand
    size_t mules_ws_doubles = 0;
    // ---- implicit path (pbe_implicit.cu) ------------------------------------------------------------
    // every cell's faces as [imp_W][ncells] slots, ascending face order (imp_W = maxf rounded up to 4, 6 or 8; 0: not
    // built / wider cells, which walk the CSR): other as csr_other, -1 = padding; face as csr_face
    int imp_W = -1;
    int32_t *d_imp_other = nullptr, *d_imp_face = nullptr;
    double *imp_ws = nullptr; // solver vectors r, r0, p, v, t [nfields][ncells] + per-cell rows, grown on demand
    size_t imp_ws_doubles = 0;
    void *imp_small = nullptr; // per-field scalars and block partials
    size_t imp_small_bytes = 0;
    void *imp_pinned = nullptr; // host mirror of the per-field scalars
    size_t imp_pinned_bytes = 0;
    cudaEvent_t imp_ev[2] = {nullptr, nullptr};
    // decomposed solve (pbe_mesh_set_halo_plan): neighbour q = rank halo_peer[q] shares the coupled faces
    // [halo_off[q], halo_off[q] + halo_cnt[q]) (indices into the coupled range); d_send_cells [ncoupled]: this rank's cell
    // on every coupled face
    std::vector<int> halo_peer;
    std::vector<int64_t> halo_off, halo_cnt;
    int32_t *d_send_cells = nullptr;
    int64_t *d_halo_qoff = nullptr, *d_halo_qcnt = nullptr; // per coupled face: its neighbour's first face and face count
    int64_t global_ncells = 0;
    double *imp_halo = nullptr; // send [nfields][ncoupled] | recv [nfields][ncoupled]
    size_t imp_halo_doubles = 0;
};
