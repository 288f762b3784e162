// api.cu -- the four kernel entry points of include/b200lk.h.
//
// b200lk_update_partials turns the host's post-order TraversalInfo list into
//   (1) one tables launch (P-factor tables for every child, on device),
//   (2) one update launch per LEVEL: nodes whose internal children are all at lower levels are
//       independent and share a launch (grid.y = nodes of the level).
// Only the node/child descriptors (16 + 32 bytes each, incl. branch lengths) cross PCIe.
#include <algorithm>

#include "common.cuh"

static int check_ready(b200lk_ctx *c) {
    if (!c) return B200LK_ERR_ARG;
    if (!c->model_set || !c->tips_set) {
        snprintf(c->err, sizeof(c->err), "upload_tips and set_model must precede kernel calls");
        return B200LK_ERR_STATE;
    }
    return B200LK_OK;
}

int enqueue_update(b200lk_ctx *c, const b200lk_node *nodes, int32_t nnodes, const b200lk_child *children,
                          int32_t nchildren) {
    if (nnodes == 0) return B200LK_OK;
    if (!nodes || !children || nnodes < 0 || nchildren < 0) return B200LK_ERR_ARG;
    if (c->ntot == 0) return B200LK_OK;  // no patterns (e.g. a trailing shard of a small alignment): nothing to launch
    if (nnodes > c->nodes_cap || nchildren > c->children_cap) {
        snprintf(c->err, sizeof(c->err), "traversal too long: %d nodes / %d children (cap %d / %d)", nnodes,
                 nchildren, c->nodes_cap, c->children_cap);
        return B200LK_ERR_ARG;
    }
    if (c->nonrev) return enqueue_update_nonrev(c, nodes, nnodes, children, nchildren);
    const int n = c->d.nstates, block = c->block, nslots = c->d.nslots;
    // next slice of the pinned staging ring; wait only if its previous copy is still in flight
    const int ring = c->stage_idx++ % B200_STAGE_RING;
    if (c->stage_used[ring]) CUDA_TRY(c, cudaEventSynchronize(c->stage_ev[ring]));
    c->h_nodes = c->h_nodes_ring + (size_t)ring * c->nodes_cap;
    c->h_children = c->h_children_ring + (size_t)ring * c->children_cap;
    c->h_prog = (char *)c->h_prog_ring + (size_t)ring * c->prog_stride;
    // level of each node = 1 + max level of the children produced earlier in this same list
    // (slots may be recycled inside one list in memory-saving mode: a write also waits for earlier
    // readers and writers of the same slot, so batching never reorders a hazard)
    std::vector<int> level_of_slot(nslots, -1), reader_level(nslots, -1), level(nnodes, 0);
    int nlevels = 0;
    for (int i = 0; i < nnodes; i++) {
        const b200lk_node &nd = nodes[i];
        if (nd.out_slot < 0 || nd.out_slot >= nslots || nd.nchild < 2 || nd.child_begin < 0 ||
            nd.child_begin + nd.nchild > nchildren) {
            snprintf(c->err, sizeof(c->err), "bad traversal entry %d", i);
            return B200LK_ERR_ARG;
        }
        int lv = 0;
        for (int k = 0; k < nd.nchild; k++) {
            const b200lk_child &ch = children[nd.child_begin + k];
            if (ch.kind == B200LK_CHILD_LEAF) {
                if (ch.id < 0 || ch.id >= c->d.ntaxa) return B200LK_ERR_ARG;
            } else {
                if (ch.id < 0 || ch.id >= nslots || ch.id == nd.out_slot) return B200LK_ERR_ARG;
                lv = std::max(lv, level_of_slot[ch.id] + 1);
            }
        }
        lv = std::max(lv, std::max(reader_level[nd.out_slot], level_of_slot[nd.out_slot]) + 1);
        for (int k = 0; k < nd.nchild; k++) {
            const b200lk_child &ch = children[nd.child_begin + k];
            if (ch.kind != B200LK_CHILD_LEAF) reader_level[ch.id] = std::max(reader_level[ch.id], lv);
        }
        level[i] = lv;
        level_of_slot[nd.out_slot] = lv;
        nlevels = std::max(nlevels, lv + 1);
    }
    // children: table offsets; the reference puts a leaf child first ("left") when exactly one child
    // of a bifurcating node is a leaf (tree/phylokernelnew.h:1395-1405)
    size_t off = 0;
    for (int j = 0; j < nchildren; j++) {
        DevChild &dc = c->h_children[j];
        dc.kind = children[j].kind;
        dc.id = children[j].id;
        dc.length = children[j].length;
        dc.echild_off = (int64_t)off;
        off += (size_t)block * n + block;  // echild, then expv
        dc.frag_off = -1;
        dc.aux_off = -1;
        dc.vtt_off = dc.vpair_off = -1;
        dc.src = j;
        dc.len_slot = c->cur_len_slots ? c->cur_len_slots[j] : -1;
        if (c->frag_cat) {
            dc.frag_off = (int64_t)off;
            off += (size_t)c->ncat_mix * c->frag_cat;
        }
        if (dc.kind == B200LK_CHILD_LEAF) {
            dc.leaf_off = (int64_t)off;
            off += (size_t)c->d.n_state_codes * block;
        } else {
            dc.leaf_off = -1;
        }
    }
    if (off > c->tables_cap) return B200LK_ERR_NOMEM;
    if (c->tile == 32) {
        // 4-state path (s4.cu): every run of bifurcating nodes is ONE pattern-stationary launch;
        // a multifurcating node is its own launch in between (stream order keeps the dependencies)
        CUDA_TRY(c, cudaMemcpyAsync(c->d_children, c->h_children, sizeof(DevChild) * nchildren,
                                    cudaMemcpyHostToDevice, c->stream));
        int rc = s4_before_update(c, nodes, nnodes, children);  // virtual tip-tip vectors this list reads or leaves alone
        if (rc) return rc;
        rc = launch_tables(c, nchildren);
        if (rc) return rc;
        bool any_multi = false;
        for (int i = 0; i < nnodes; i++) {
            DevNode &dn = c->h_nodes[i];
            dn.out_slot = nodes[i].out_slot;
            dn.child_begin = nodes[i].child_begin;
            dn.nchild = nodes[i].nchild;
            dn.kind = nodes[i].nchild > 2 ? 3 : 2;
            any_multi |= nodes[i].nchild > 2;
        }
        if (any_multi)
            CUDA_TRY(c, cudaMemcpyAsync(c->d_nodes, c->h_nodes, sizeof(DevNode) * nnodes, cudaMemcpyHostToDevice,
                                        c->stream));
        int prog_used = 0, seq_used = 0;
        c->s4_tt_used = 0;
        c->s4_ttdesc_used = 0;
        // how many nodes of THIS list read each node's output (before its slot is rewritten): a tip-tip vector may stay
        // virtual only if its single reader takes it from registers or the stack (s4.cu)
        std::vector<int> last_writer(nslots, -1), nreaders(nnodes, 0);
        for (int i = 0; i < nnodes; i++) {
            for (int k = 0; k < nodes[i].nchild; k++) {
                const b200lk_child &ch = children[nodes[i].child_begin + k];
                if (ch.kind != B200LK_CHILD_LEAF && last_writer[ch.id] >= 0) nreaders[last_writer[ch.id]]++;
            }
            last_writer[nodes[i].out_slot] = i;
        }
        c->cur_nodes = nodes;
        c->cur_nreaders = nreaders.data();
        for (int i = 0; i < nnodes;) {
            if (nodes[i].nchild > 2) {
                rc = launch_partial_multi(c, i);
                if (rc) return rc;
                i++;
                continue;
            }
            int j = i;
            while (j < nnodes && nodes[j].nchild == 2) j++;
            rc = launch_s4_traverse(c, nodes, i, j - i, &prog_used, &seq_used);
            if (rc) return rc;
            i = j;
        }
        c->cur_nodes = nullptr;
        c->cur_nreaders = nullptr;
        CUDA_TRY(c, cudaEventRecord(c->stage_ev[ring], c->stream));
        c->stage_used[ring] = true;
        c->n_updates += (int64_t)nnodes * c->ntot;
        c->theta_computed = false;
        return B200LK_OK;
    }
    // level-sorted node array; inside a level: tip-tip nodes (20-state path: answered from a per-node table over
    // state pairs, dense.cu), then the other bifurcating nodes, then multifurcating ones
    std::vector<int> kind_of(nnodes, 3);
    for (int i = 0; i < nnodes; i++) {
        if (nodes[i].nchild > 2) continue;
        DevChild &a = c->h_children[nodes[i].child_begin], &b = c->h_children[nodes[i].child_begin + 1];
        if (a.kind != B200LK_CHILD_LEAF && b.kind == B200LK_CHILD_LEAF) std::swap(a, b);  // a single leaf goes first
        kind_of[i] = (a.kind == B200LK_CHILD_LEAF) ? (b.kind == B200LK_CHILD_LEAF ? 0 : 1) : 2;
    }
    // 20-state path: a tip-tip node answered from its state-pair table whose ONLY reader in this list is a bifurcating
    // node of the tensor kernel stays VIRTUAL -- it is not gathered into its slot; the reader addresses the table itself
    // (dense20_rp_kernel), and anybody else who wants the slot later gets it gathered first (s4_realize).  A third of the
    // vectors of a full traversal are such nodes; their gather launches were 12 % of a cfg3 evaluation.
    const bool tt_ok = c->dense20 && !c->batch && b200_env().s4_tt &&
                       (b200_env().s4_tt == 2 || (int64_t)c->d.n_state_codes * c->d.n_state_codes * 4 <= c->ntot);
    const bool virt_ok = tt_ok && c->virt && b200_env().s4_virtual && b200_env().dense_pipe == 2 && !c->site;
    std::vector<int> nreaders(nnodes, 0), reader(nnodes, -1);
    std::vector<char> vnode(nnodes, 0);
    if (c->virt) {
        int rb = s4_before_update(c, nodes, nnodes, children);  // the tables are about to be rebuilt
        if (rb) return rb;
    }
    if (virt_ok) {
        std::vector<int> last_writer(nslots, -1);
        for (int i = 0; i < nnodes; i++) {
            for (int k = 0; k < nodes[i].nchild; k++) {
                const b200lk_child &ch = children[nodes[i].child_begin + k];
                if (ch.kind != B200LK_CHILD_LEAF && last_writer[ch.id] >= 0) {
                    nreaders[last_writer[ch.id]]++;
                    reader[last_writer[ch.id]] = i;
                }
            }
            last_writer[nodes[i].out_slot] = i;
        }
        for (int i = 0; i < nnodes; i++)
            vnode[i] = kind_of[i] == 0 && nreaders[i] == 1 && nodes[reader[i]].nchild == 2;
    }
    std::vector<std::vector<int>> levels(nlevels);
    int pos = 0;
    size_t tt_off = 0;
    const size_t tt_doubles = (size_t)c->d.n_state_codes * c->d.n_state_codes * block;
    std::vector<S4Virt> vdesc;          // the virtual nodes of this list ...
    std::vector<int64_t> vpair_off;     // ... and where their pair indices go
    for (int i = 0; i < nnodes && virt_ok; i++) {
        if (!vnode[i]) continue;
        if (tt_off + tt_doubles > (size_t)INT32_MAX) { vnode[i] = 0; continue; }  // descriptors carry 32-bit offsets
        const DevChild &a = c->h_children[nodes[i].child_begin], &b = c->h_children[nodes[i].child_begin + 1];
        const S4Virt d = {nodes[i].out_slot, (int32_t)tt_off, a.id, b.id};
        // the reader's entry for this child
        const b200lk_node &rn = nodes[reader[i]];
        for (int k = 0; k < 2; k++) {
            DevChild &rc = c->h_children[rn.child_begin + k];
            if (rc.kind != B200LK_CHILD_LEAF && rc.id == nodes[i].out_slot) {
                rc.vtt_off = (int64_t)tt_off;
                rc.vpair_off = (int64_t)vdesc.size() * c->ntot_pad;
            }
        }
        c->h_children[nodes[i].child_begin].aux_off = (int64_t)tt_off;  // tt_tables_kernel builds the table as for any tip-tip node
        vpair_off.push_back((int64_t)vdesc.size() * c->ntot_pad);
        vdesc.push_back(d);
        tt_off += tt_doubles;
    }
    for (int lv = 0; lv < nlevels; lv++) {
        // inside a level: tip-tip nodes, bifurcating nodes without a virtual child (lean tensor kernel), bifurcating nodes
        // with one (the build that reads state-pair tables), multifurcating nodes
        for (int pass = 0; pass < 4; pass++)
            for (int i = 0; i < nnodes; i++) {
                if (level[i] != lv || vnode[i]) continue;
                int want = kind_of[i] == 0 ? 0 : (kind_of[i] == 3 ? 3 : 1);
                if (want == 1 && b200_env().dense_split_virt)
                    for (int k = 0; k < 2; k++)
                        if (c->h_children[nodes[i].child_begin + k].vtt_off >= 0) want = 2;
                if (want != pass) continue;
                DevNode &dn = c->h_nodes[pos];
                dn.out_slot = nodes[i].out_slot;
                dn.child_begin = nodes[i].child_begin;
                dn.nchild = nodes[i].nchild;
                dn.kind = kind_of[i];
                // state-pair table of a tip-tip node: worth building when there are clearly more patterns than pairs
                if (dn.kind == 0 && tt_ok) {
                    c->h_children[dn.child_begin].aux_off = (int64_t)tt_off;
                    tt_off += tt_doubles;
                }
                levels[lv].push_back(pos);
                pos++;
            }
    }
    // virtual nodes: behind every level in the device node array (tt_tables_kernel visits them, no update launch does)
    for (int i = 0; i < nnodes; i++) {
        if (!vnode[i]) continue;
        DevNode &dn = c->h_nodes[pos++];
        dn.out_slot = nodes[i].out_slot;
        dn.child_begin = nodes[i].child_begin;
        dn.nchild = 2;
        dn.kind = 0;
    }
    // bookkeeping in list order: the last writer of a slot decides whether it is virtual
    if (c->virt)
        for (int i = 0, k = 0; i < nnodes; i++) {
            const int slot = nodes[i].out_slot;
            if (c->virt[slot] && !vnode[i]) c->n_virtual--;
            if (!c->virt[slot] && vnode[i]) c->n_virtual++;
            c->virt[slot] = vnode[i];
            if (vnode[i]) c->virt_desc[slot] = vdesc[k++];
        }
    if (tt_off > c->tt_cap) {
        if (c->tt_tables) CUDA_TRY(c, cudaFree(c->tt_tables));
        c->tt_tables = nullptr;
        c->tt_cap = 0;
        CUDA_TRY(c, cudaMalloc(&c->tt_tables, sizeof(double) * tt_off));
        c->tt_cap = tt_off;
    }
    CUDA_TRY(c, cudaMemcpyAsync(c->d_children, c->h_children, sizeof(DevChild) * nchildren, cudaMemcpyHostToDevice,
                                c->stream));
    CUDA_TRY(c, cudaMemcpyAsync(c->d_nodes, c->h_nodes, sizeof(DevNode) * nnodes, cudaMemcpyHostToDevice, c->stream));
    CUDA_TRY(c, cudaEventRecord(c->stage_ev[ring], c->stream));
    c->stage_used[ring] = true;
    if (c->site) {  // site-specific models: no P-factor tables, per-pattern eigen-systems (site.cu)
        int rs = launch_site_partials(c, levels);
        if (rs) return rs;
        c->n_updates += (int64_t)nnodes * c->ntot;
        c->theta_computed = false;
        return B200LK_OK;
    }
    if (!c->batch) {  // level-batched kernels fetch every internal child from HBM and write every vector once
        const int64_t vec_bytes = 8 * (int64_t)c->block + c->scale_per_ptn;
        int64_t vecs = nnodes - 2 * (int64_t)vdesc.size(), tips = 2 * (int64_t)vdesc.size();  // a virtual node: no write, no read, two tip rows at its reader
        for (int j = 0; j < nchildren; j++) (children[j].kind == B200LK_CHILD_LEAF ? tips : vecs)++;
        c->n_min_bytes += c->ntot * (vec_bytes * vecs + tips);
    }
    int rc = launch_tables(c, nchildren);
    if (rc) return rc;
    if (tt_off > 0) {
        rc = launch_tt_tables(c, nnodes);
        if (rc) return rc;
    }
    if (!vdesc.empty()) {
        rc = launch_tt_pairs(c, vdesc.data(), (int)vdesc.size(), vpair_off.data());
        if (rc) return rc;
    }
    rc = launch_partials(c, levels);
    if (rc) return rc;
    c->n_updates += (int64_t)nnodes * c->ntot;
    c->theta_computed = false;
    return B200LK_OK;
}

int check_ends(b200lk_ctx *c, b200lk_end dad, b200lk_end node) {
    if (c->nonrev) return nr_check_ends(c, dad, node);
    if (node.kind != B200LK_CHILD_INTERNAL || node.id < 0 || node.id >= c->d.nslots) return B200LK_ERR_ARG;
    if (dad.kind == B200LK_CHILD_LEAF) return (dad.id >= 0 && dad.id < c->d.ntaxa) ? B200LK_OK : B200LK_ERR_ARG;
    if (dad.kind != B200LK_CHILD_INTERNAL || dad.id < 0 || dad.id >= c->d.nslots || dad.id == node.id)
        return B200LK_ERR_ARG;
    return B200LK_OK;
}

// copy `count` result doubles to the pinned host buffer and wait: the only D2H of an evaluation
int read_results(b200lk_ctx *c, int first, int count) {
    CUDA_TRY(c, cudaMemcpyAsync(c->h_result + first, c->d_result + first, sizeof(double) * count,
                                cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    return B200LK_OK;
}

int finish_lnl(b200lk_ctx *c, double *lnl, bool from_branch) {
    int rc = read_results(c, 0, 1);
    if (rc) return rc;
    int warn = 0;
    if (!isfinite(c->h_result[0])) {  // tree/phylokernelnew.h:3077-3094
        warn = B200LK_WARN_LNL_NONFINITE;
        rc = launch_repair(c);
        if (rc) return rc;
        if (from_branch && (c->robust_keep < 1.0 || c->robust_median)) {  // the robust score is taken from the repaired per-pattern values (:3096)
            rc = launch_robust(c);
            if (rc) return rc;
        }
        rc = read_results(c, 0, 1);
        if (rc) return rc;
    }
    *lnl = c->h_result[0];
    return warn;
}

int finish_derv(b200lk_ctx *c, double *df, double *ddf) {
    int rc = read_results(c, 1, 2);
    if (rc) return rc;
    double a = c->h_result[1], b = c->h_result[2];
    int warn = 0;
    if (!isfinite(a)) {  // tree/phylokernelnew.h:2573-2578, 2629-2632
        warn = B200LK_WARN_DERV_NONFINITE;
        if (c->d.safe_numeric) a = b = 0.0;
    }
    if (df) *df = a;
    if (ddf) *ddf = b;
    return warn | derv_info_bits(c);
}

extern "C" int b200lk_update_partials(b200lk_ctx *c, const b200lk_node *nodes, int32_t nnodes,
                                      const b200lk_child *children, int32_t nchildren) {
    SHARD_FORWARD(c, shard_update_partials(c, nodes, nnodes, children, nchildren));
    int rc = check_ready(c);
    if (rc) return rc;
    CUDA_TRY(c, cudaSetDevice(c->device));
    rc = enqueue_update(c, nodes, nnodes, children, nchildren);
    if (rc) return rc;
    return B200LK_OK;  // asynchronous: later calls on this context are stream-ordered after it
}

extern "C" int b200lk_branch_lnl(b200lk_ctx *c, b200lk_end dad, b200lk_end node, double length, double *lnl) {
    SHARD_FORWARD(c, shard_branch_lnl(c, dad, node, length, lnl));
    int rc = check_ready(c);
    if (rc) return rc;
    if (!lnl || check_ends(c, dad, node)) return B200LK_ERR_ARG;
    CUDA_TRY(c, cudaSetDevice(c->device));
    if (c->d.nptn == 0) { *lnl = 0.0; return B200LK_OK; }
    rc = launch_branch_lnl(c, dad, node, length);
    if (rc) return rc;
    return finish_lnl(c, lnl, true);
}

extern "C" int b200lk_build_theta(b200lk_ctx *c, b200lk_end dad, b200lk_end node) {
    SHARD_FORWARD(c, shard_build_theta(c, dad, node));
    int rc = check_ready(c);
    if (rc) return rc;
    if (check_ends(c, dad, node)) return B200LK_ERR_ARG;
    CUDA_TRY(c, cudaSetDevice(c->device));
    if (c->nonrev) {
        nr_record_ends(c, dad, node);
    } else if (c->ntot > 0) {
        rc = launch_theta(c, dad, node);
        if (rc) return rc;
    }
    c->theta_computed = true;
    return B200LK_OK;
}

extern "C" int b200lk_derv(b200lk_ctx *c, double length, double *df, double *ddf) {
    SHARD_FORWARD(c, shard_derv(c, length, df, ddf));
    int rc = check_ready(c);
    if (rc) return rc;
    if (!c->theta_computed) return B200LK_ERR_STATE;
    CUDA_TRY(c, cudaSetDevice(c->device));
    if (c->d.nptn == 0) { if (df) *df = 0; if (ddf) *ddf = 0; return B200LK_OK; }
    rc = launch_derv(c, length);
    if (rc) return rc;
    return finish_derv(c, df, ddf);
}

extern "C" int b200lk_lnl_from_buffer(b200lk_ctx *c, double length, double *lnl) {
    SHARD_FORWARD(c, shard_lnl_from_buffer(c, length, lnl));
    int rc = check_ready(c);
    if (rc) return rc;
    if (!lnl) return B200LK_ERR_ARG;
    if (!c->theta_computed) return B200LK_ERR_STATE;
    CUDA_TRY(c, cudaSetDevice(c->device));
    if (c->d.nptn == 0) { *lnl = 0.0; return B200LK_OK; }
    rc = launch_from_buffer(c, length);
    if (rc) return rc;
    return finish_lnl(c, lnl, false);
}

extern "C" int b200lk_optimize_branch(b200lk_ctx *c, b200lk_end dad, b200lk_end node, const b200lk_newton *nr,
                                      double *optx, double *d2l, int32_t *nsteps) {
    SHARD_FORWARD(c, shard_optimize_branch(c, dad, node, nr, optx, d2l, nsteps));
    int rc = check_ready(c);
    if (rc) return rc;
    if (!nr || !optx || check_ends(c, dad, node) || !(nr->min_len <= nr->max_len) || nr->max_steps < 1)
        return B200LK_ERR_ARG;
    if (c->nonrev) return B200LK_ERR_UNSUPPORTED;  // no theta buffer: the host iteration drives b200lk_derv
    CUDA_TRY(c, cudaSetDevice(c->device));
    if (c->d.nptn == 0) {
        *optx = nr->guess;
        if (d2l) *d2l = 0.0;
        if (nsteps) *nsteps = 0;
        return B200LK_OK;
    }
    const b200lk_end ends[2] = {dad, node};
    rc = launch_newton(c, nr, ends);  // theta for the branch, then the iteration (one launch on the T32 layout)
    if (rc == B200LK_OK || rc == B200LK_ERR_UNSUPPORTED) c->theta_computed = true;  // refused: theta is built all the same
    if (rc) return rc;
    rc = read_results(c, 4, 4);  // optx, d2l, evaluations, status: the only D2H of the whole iteration
    if (rc) return rc;
    *optx = c->h_result[4];
    if (d2l) *d2l = c->h_result[5];
    if (nsteps) *nsteps = (int32_t)c->h_result[6];
    return (c->h_result[7] != 0.0 ? B200LK_WARN_DERV_NONFINITE : B200LK_OK) | derv_info_bits(c);
}

extern "C" int b200lk_evaluate_async(b200lk_ctx *c, const b200lk_node *nodes, int32_t nnodes,
                                     const b200lk_child *children, int32_t nchildren, b200lk_end dad,
                                     b200lk_end node, double length, int want_lnl, int want_derv) {
    SHARD_FORWARD(c, shard_evaluate_async(c, nodes, nnodes, children, nchildren, dad, node, length, want_lnl, want_derv));
    int rc = check_ready(c);
    if (rc) return rc;
    if (check_ends(c, dad, node)) return B200LK_ERR_ARG;
    CUDA_TRY(c, cudaSetDevice(c->device));
    rc = enqueue_update(c, nodes, nnodes, children, nchildren);
    if (rc) return rc;
    if (c->d.nptn == 0) {  // an empty shard still takes part in the exchange, or its peers would wait for ever
        CUDA_TRY(c, cudaMemsetAsync(c->d_result, 0, sizeof(double) * 3, c->stream));
        return launch_peer_exchange(c, 3);
    }
    if (want_lnl && want_derv && can_fuse_tail(c)) {  // one launch for lnL + theta + df/ddf (branch.cu)
        rc = launch_eval_tail(c, dad, node, length);
        if (rc) return rc;
        c->theta_computed = true;
        return B200LK_OK;
    }
    if (want_lnl) {
        rc = launch_branch_lnl(c, dad, node, length);
        if (rc) return rc;
    }
    if (want_derv) {
        rc = launch_theta(c, dad, node);
        if (rc) return rc;
        c->theta_computed = true;
        rc = launch_derv(c, length);
        if (rc) return rc;
    }
    return launch_peer_exchange(c, 3);  // pattern shards in other processes (b200lk_peer_connect): sums over all ranks
}

extern "C" int b200lk_read_results(b200lk_ctx *c, double *lnl, double *df, double *ddf) {
    SHARD_FORWARD(c, shard_read_results(c, lnl, df, ddf));
    if (!c) return B200LK_ERR_ARG;
    CUDA_TRY(c, cudaSetDevice(c->device));
    // one synchronising read for everything
    int warn = 0;
    int rc = read_results(c, 0, 3);
    if (rc) return rc;
    if (lnl && c->peer_world > 1 && !isfinite(c->h_result[0])) {
        // the TOTAL is non-finite, on every rank alike: each repairs its own patterns (tree/phylokernelnew.h:3077-3094),
        // then the repaired sums are exchanged again
        warn |= B200LK_WARN_LNL_NONFINITE;
        if (c->d.nptn > 0) {
            rc = launch_repair(c);
            if (rc) return rc;
        }
        rc = launch_peer_exchange(c, 1);
        if (rc) return rc;
        rc = read_results(c, 0, 1);
        if (rc) return rc;
        *lnl = c->h_result[0];
    } else if (lnl) {
        if (!isfinite(c->h_result[0]) && c->d.nptn > 0) {
            rc = finish_lnl(c, lnl, true);
            if (rc < 0) return rc;
            warn |= rc;
        } else {
            *lnl = c->h_result[0];
        }
    }
    if (df || ddf) {
        double a = c->h_result[1], b = c->h_result[2];
        if (!isfinite(a)) {
            warn |= B200LK_WARN_DERV_NONFINITE;
            if (c->d.safe_numeric) a = b = 0.0;
        }
        if (df) *df = a;
        if (ddf) *ddf = b;
        warn |= derv_info_bits(c);
    }
    return warn;
}

extern "C" int b200lk_evaluate(b200lk_ctx *c, const b200lk_node *nodes, int32_t nnodes, const b200lk_child *children,
                               int32_t nchildren, b200lk_end dad, b200lk_end node, double length, double *lnl,
                               double *df, double *ddf) {
    int rc = b200lk_evaluate_async(c, nodes, nnodes, children, nchildren, dad, node, length, lnl != nullptr,
                                   df != nullptr || ddf != nullptr);
    if (rc) return rc;
    return b200lk_read_results(c, lnl, df, ddf);
}

// params->robust_phy_keep / robust_median as computeLikelihoodBranch reads them (tree/phylokernelnew.h:3096-3135)
extern "C" int b200lk_set_robust(b200lk_ctx *c, double keep, int32_t median) {
    SHARD_FORWARD(c, shard_set_robust(c, keep, median));
    if (!c || !(keep > 0.0) || keep > 1.0) return B200LK_ERR_ARG;
    if ((keep < 1.0 || median) && (c->nonrev || c->site)) return B200LK_ERR_UNSUPPORTED;
    c->robust_keep = keep;
    c->robust_median = median != 0;
    return B200LK_OK;
}

// ---- +ASC flavour, pattern weights --------------------------------------------------------------------------------
extern "C" int b200lk_set_asc(b200lk_ctx *c, int32_t type) {
    SHARD_REFUSE(c, "+ASC");
    if (!c || (type != B200LK_ASC_LEWIS && type != B200LK_ASC_HOLDER)) return B200LK_ERR_ARG;
    if (type == B200LK_ASC_HOLDER) {
        if (c->d.n_unobserved == 0 || c->d.n_unobserved != (int64_t)c->d.nstates * c->d.nptn) {
            snprintf(c->err, sizeof(c->err), "Holder's +ASC needs n_unobserved == nstates * nptn");
            return B200LK_ERR_ARG;
        }
        if (c->site || c->nonrev) return B200LK_ERR_UNSUPPORTED;
        CUDA_TRY(c, cudaSetDevice(c->device));
        if (!c->const_df) {
            CUDA_TRY(c, cudaMalloc(&c->const_df, sizeof(double) * c->d.n_unobserved));
            CUDA_TRY(c, cudaMalloc(&c->const_ddf, sizeof(double) * c->d.n_unobserved));
        }
    }
    c->asc_holder = type == B200LK_ASC_HOLDER;
    return B200LK_OK;
}

extern "C" int b200lk_set_ptn_freq(b200lk_ctx *c, const double *ptn_freq) {
    SHARD_FORWARD(c, shard_set_ptn_freq(c, ptn_freq));
    if (!c || !ptn_freq) return B200LK_ERR_ARG;
    if (!c->tips_set) return B200LK_ERR_STATE;
    CUDA_TRY(c, cudaSetDevice(c->device));
    // nsite stays aln->getNSite(): the EM step of the heterotachy model only re-weights the patterns
    CUDA_TRY(c, cudaMemcpyAsync(c->ptn_freq, ptn_freq, sizeof(double) * c->d.nptn, cudaMemcpyHostToDevice, c->stream));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    return B200LK_OK;
}

// ---- heterotachy -----------------------------------------------------------------------------------------------------
static int check_mixlen(b200lk_ctx *c) {
    int rc = check_ready(c);
    if (rc) return rc;
    if (c->site || c->nonrev || c->batch) {
        snprintf(c->err, sizeof(c->err), "mixed branch lengths: reversible, non-site-specific contexts only");
        return B200LK_ERR_UNSUPPORTED;
    }
    return B200LK_OK;
}

// class lengths of the branch of a branch-level call -> device; the launchers read c->cur_branch_mixlen
static int stage_branch_mixlen(b200lk_ctx *c, const double *len) {
    if (!c->d_branch_mixlen) CUDA_TRY(c, cudaMalloc(&c->d_branch_mixlen, sizeof(double) * c->d.ncat));
    CUDA_TRY(c, cudaMemcpyAsync(c->d_branch_mixlen, len, sizeof(double) * c->d.ncat, cudaMemcpyHostToDevice, c->stream));
    c->cur_branch_mixlen = c->d_branch_mixlen;
    return B200LK_OK;
}

extern "C" int b200lk_update_partials_mixlen(b200lk_ctx *c, const b200lk_node *nodes, int32_t nnodes,
                                             const b200lk_child *children, int32_t nchildren, const double *child_len) {
    SHARD_REFUSE(c, "heterotachy (mixed branch lengths)");
    int rc = check_mixlen(c);
    if (rc) return rc;
    if (!child_len) return B200LK_ERR_ARG;
    if (nnodes == 0) return B200LK_OK;
    if (nchildren < 0 || nchildren > c->children_cap) return B200LK_ERR_ARG;
    CUDA_TRY(c, cudaSetDevice(c->device));
    if (!c->d_child_mixlen)
        CUDA_TRY(c, cudaMalloc(&c->d_child_mixlen, sizeof(double) * (size_t)c->children_cap * c->d.ncat));
    // pageable source: the copy is staged before the call returns, so the caller may reuse child_len
    CUDA_TRY(c, cudaMemcpyAsync(c->d_child_mixlen, child_len, sizeof(double) * (size_t)nchildren * c->d.ncat,
                                cudaMemcpyHostToDevice, c->stream));
    c->h_child_mixlen = child_len;
    rc = enqueue_update(c, nodes, nnodes, children, nchildren);
    c->h_child_mixlen = nullptr;
    return rc;
}

extern "C" int b200lk_branch_lnl_mixlen(b200lk_ctx *c, b200lk_end dad, b200lk_end node, const double *len, double *lnl) {
    SHARD_REFUSE(c, "heterotachy (mixed branch lengths)");
    int rc = check_mixlen(c);
    if (rc) return rc;
    if (!len || !lnl || check_ends(c, dad, node)) return B200LK_ERR_ARG;
    CUDA_TRY(c, cudaSetDevice(c->device));
    if (c->d.nptn == 0) { *lnl = 0.0; return B200LK_OK; }
    rc = stage_branch_mixlen(c, len);
    if (rc) return rc;
    rc = launch_branch_lnl(c, dad, node, len[0]);
    c->cur_branch_mixlen = nullptr;
    if (rc) return rc;
    return finish_lnl(c, lnl, true);
}

extern "C" int b200lk_lnl_from_buffer_mixlen(b200lk_ctx *c, const double *len, double *lnl) {
    SHARD_REFUSE(c, "heterotachy (mixed branch lengths)");
    int rc = check_mixlen(c);
    if (rc) return rc;
    if (!len || !lnl) return B200LK_ERR_ARG;
    if (!c->theta_computed) return B200LK_ERR_STATE;
    CUDA_TRY(c, cudaSetDevice(c->device));
    if (c->d.nptn == 0) { *lnl = 0.0; return B200LK_OK; }
    rc = stage_branch_mixlen(c, len);
    if (rc) return rc;
    rc = launch_from_buffer(c, len[0]);
    c->cur_branch_mixlen = nullptr;
    if (rc) return rc;
    return finish_lnl(c, lnl, false);
}

extern "C" int b200lk_derv_mixlen(b200lk_ctx *c, const double *len, double *df, double *ddf, double *lnl) {
    SHARD_REFUSE(c, "heterotachy (mixed branch lengths)");
    int rc = check_mixlen(c);
    if (rc) return rc;
    if (!len || !df || !ddf) return B200LK_ERR_ARG;
    if (!c->theta_computed) return B200LK_ERR_STATE;
    const int ncat = c->d.ncat, K = 1 + ncat + ncat * ncat;
    if (ncat > 8 || c->d.n_unobserved > 0) {  // the reference asserts on +ASC here (tree/phylokernelnew.h:2449)
        snprintf(c->err, sizeof(c->err), "derv_mixlen: at most 8 classes, no +ASC");
        return B200LK_ERR_UNSUPPORTED;
    }
    CUDA_TRY(c, cudaSetDevice(c->device));
    if (c->d.nptn == 0) {
        for (int i = 0; i < ncat; i++) df[i] = 0.0;
        for (int i = 0; i < ncat * ncat; i++) ddf[i] = 0.0;
        if (lnl) *lnl = 0.0;
        return B200LK_OK;
    }
    if (!c->mix_red) {
        CUDA_TRY(c, cudaMalloc(&c->mix_red, sizeof(double) * (size_t)(1 + 8 + 64) * c->red_blocks));
        CUDA_TRY(c, cudaMalloc(&c->mix_result, sizeof(double) * (1 + 8 + 64)));
    }
    rc = stage_branch_mixlen(c, len);
    if (rc) return rc;
    rc = launch_derv_mixlen(c, K);
    c->cur_branch_mixlen = nullptr;
    if (rc) return rc;
    double out[1 + 8 + 64];
    CUDA_TRY(c, cudaMemcpyAsync(out, c->mix_result, sizeof(double) * K, cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    if (lnl) *lnl = out[0];
    for (int i = 0; i < ncat; i++) df[i] = out[1 + i];
    for (int i = 0; i < ncat * ncat; i++) ddf[i] = out[1 + ncat + i];
    for (int i = 0; i < ncat; i++)
        if (!isfinite(df[i])) return B200LK_WARN_DERV_NONFINITE;  // ASSERT(isfinite(df[i])) at :2560
    return B200LK_OK;
}

extern "C" int b200lk_derv_mixlen_class(b200lk_ctx *c, int32_t cur_mixlen, double len, double *df, double *ddf) {
    SHARD_REFUSE(c, "heterotachy (mixed branch lengths)");
    int rc = check_mixlen(c);
    if (rc) return rc;
    if (cur_mixlen < 0 || cur_mixlen >= c->d.ncat) return B200LK_ERR_ARG;
    if (!c->theta_computed) return B200LK_ERR_STATE;
    if (c->asc_holder || (!c->d.fused_mix_rate && c->d.nmix > 1)) {  // ASSERT(!ASC_Holder) at :3432; EM needs fused rates
        snprintf(c->err, sizeof(c->err), "derv_mixlen_class: fused_mix_rate or nmix == 1, no Holder +ASC");
        return B200LK_ERR_UNSUPPORTED;
    }
    CUDA_TRY(c, cudaSetDevice(c->device));
    if (c->d.nptn == 0) { if (df) *df = 0; if (ddf) *ddf = 0; return B200LK_OK; }
    rc = launch_derv_mixlen_class(c, cur_mixlen, len);
    if (rc) return rc;
    rc = read_results(c, 1, 2);
    if (rc) return rc;
    double a = c->h_result[1], b = c->h_result[2];
    int warn = 0;
    if (!isfinite(a)) {  // :3569-3572: warn and zero, in either scaling mode
        warn = B200LK_WARN_DERV_NONFINITE;
        a = b = 0.0;
    }
    if (df) *df = a;
    if (ddf) *ddf = b;
    return warn;
}
