// tree.cpp -- the search tree of AlphaGo/mcts.py (TreeNode, :11-92; MCTS._playout, :137-169) as host-side native
// code, driven in waves by rocalphago_b200/mcts.py.
//
// The reference keeps one Python object per child and walks one playout at a time.  Here a node holds its
// children's statistics as arrays (priors P, values Q, bonuses u, visit counts n, virtual visits of the wave in
// flight), `count` playouts advance one ply per call, and the only thing the caller does between calls is what
// needs the GPU: evaluate the positions that stand on unexpanded nodes (one policy batch), play the chosen moves
// (one rag_do_move launch), score the leaves (one value batch).  The arithmetic is the reference's, operation
// for operation and in its order, in doubles -- with one playout per wave the visit counts, Q and u values are
// bit-identical to a run of the reference (tests/test_gpu_mcts.py compares with values recorded from it):
//   select   child of largest Q + u, the first one on ties                       (mcts.py:40-46)
//   expand   one child per (action, prior), u starts as the prior                 (mcts.py:22, 26-38)
//   update   n += 1; Q += (v - Q) / n; u = c_puct * P * sqrt(parent.n) / (1 + n)  (mcts.py:60-80), root first,
//            the same leaf value added unflipped to every node of the walk
#include "rag_common.h"
#include <math.h>
#include <stdint.h>
#include <vector>

namespace {

struct Node {
    int parent = -1, index = 0;                 // `index`: which child of `parent` this node is
    long long visits = 0;                       // the root's own counters (a child's live in its parent's arrays)
    double q = 0.0;
    std::vector<int32_t> actions;               // flat moves x*S+y of the children; empty while a leaf
    std::vector<double> P, Q, u;
    std::vector<long long> n, virt;
    std::vector<int32_t> child;                 // node id or -1 per child, made when first walked into
    bool leaf() const { return actions.empty(); }
};

struct Step { int node, i; };

}  // namespace

struct rag_tree {
    double c_puct = 5.0, virtual_loss = 0.0;
    std::vector<Node> nodes;
    int root = 0;
    // the wave in flight
    int count = 0;
    std::vector<int> at;                        // node every playout stands on
    std::vector<char> alive;
    std::vector<std::vector<Step>> path;
};

namespace {

int child_of(rag_tree* t, int node, int i) {
    if (t->nodes[node].child[i] < 0) {
        Node c;
        c.parent = node; c.index = i;
        t->nodes.push_back(c);
        t->nodes[node].child[i] = (int)t->nodes.size() - 1;
    }
    return t->nodes[node].child[i];
}

long long visits_of(const rag_tree* t, int node) {
    const Node& nd = t->nodes[node];
    return nd.parent < 0 ? nd.visits : t->nodes[nd.parent].n[nd.index];
}

// copies the subtree under `src` of `from` into `to` (compaction when the root moves)
int copy_subtree(const std::vector<Node>& from, int src, std::vector<Node>& to, int parent, int index) {
    const int id = (int)to.size();
    to.push_back(from[src]);
    to[id].parent = parent; to[id].index = index;
    for (size_t i = 0; i < from[src].child.size(); i++)
        if (from[src].child[i] >= 0) {
            const int c = copy_subtree(from, from[src].child[i], to, id, (int)i);
            to[id].child[i] = c;
        }
    return id;
}

}  // namespace

extern "C" {

int rag_tree_create(double c_puct, rag_tree** out)
{
    if (!out) return rag::fail(RAG_EINVAL, "null argument");
    rag_tree* t = new rag_tree();
    t->c_puct = c_puct;
    t->nodes.push_back(Node());
    *out = t;
    return RAG_OK;
}

int rag_tree_destroy(rag_tree* t) { delete t; return RAG_OK; }

int rag_tree_reset(rag_tree* t)
{
    if (!t) return rag::fail(RAG_EINVAL, "null tree");
    t->nodes.clear();
    t->nodes.push_back(Node());
    t->root = 0; t->count = 0;
    return RAG_OK;
}

int rag_tree_wave_begin(rag_tree* t, int count, double virtual_loss)
{
    if (!t || count < 1) return rag::fail(RAG_EINVAL, "bad argument");
    t->count = count; t->virtual_loss = virtual_loss;
    t->at.assign(count, t->root);
    t->alive.assign(count, 1);
    t->path.assign(count, std::vector<Step>());
    return RAG_OK;
}

// rows[k]: the first live playout standing on each distinct unexpanded node, ascending.  Returns k.
int rag_tree_wave_leaves(rag_tree* t, int32_t* rows)
{
    if (!t || !rows) return rag::fail(RAG_EINVAL, "null argument");
    int k = 0;
    std::vector<int> seen;
    for (int j = 0; j < t->count; j++) {
        if (!t->alive[j] || !t->nodes[t->at[j]].leaf()) continue;
        bool dup = false;
        for (int s : seen) if (s == t->at[j]) { dup = true; break; }
        if (dup) continue;
        seen.push_back(t->at[j]);
        rows[k++] = j;
    }
    return k;
}

// expands the nodes under playouts rows[0..k): counts[i] children each, taken in order from moves / priors
int rag_tree_wave_expand(rag_tree* t, int k, const int32_t* rows, const int64_t* counts, const int64_t* moves, const float* priors)
{
    if (!t || (k > 0 && (!rows || !counts || !moves || !priors))) return rag::fail(RAG_EINVAL, "null argument");
    size_t off = 0;
    for (int r = 0; r < k; r++) {
        const int c = (int)counts[r];
        if (rows[r] < 0 || rows[r] >= t->count || c < 0) return rag::fail(RAG_EINVAL, "bad row %d", r);
        Node& nd = t->nodes[t->at[rows[r]]];
        if (c > 0 && nd.leaf()) {
            nd.actions.assign(c, 0); nd.P.assign(c, 0.0); nd.Q.assign(c, 0.0); nd.u.assign(c, 0.0);
            nd.n.assign(c, 0); nd.virt.assign(c, 0); nd.child.assign(c, -1);
            for (int i = 0; i < c; i++) {
                nd.actions[i] = (int32_t)moves[off + i];
                nd.P[i] = (double)priors[off + i];
                nd.u[i] = nd.P[i];
            }
        }
        off += (size_t)c;
    }
    return RAG_OK;
}

// every live playout takes the child of largest Q + u (minus the virtual visits of this wave) of the node it
// stands on; a playout on a node without children has reached the end of the game.  moves_out[j] = the move
// taken or -2.  Returns how many playouts are still alive.
int rag_tree_wave_step(rag_tree* t, int32_t* moves_out)
{
    if (!t || !moves_out) return rag::fail(RAG_EINVAL, "null argument");
    const double vl = t->count > 1 ? t->virtual_loss : 0.0;
    int alive = 0;
    for (int j = 0; j < t->count; j++) {
        moves_out[j] = -2;
        if (!t->alive[j]) continue;
        const int node = t->at[j];
        if (t->nodes[node].leaf()) { t->alive[j] = 0; continue; }
        int best;
        {
            Node& nd = t->nodes[node];
            best = 0;
            double best_score = 0.0;
            for (size_t i = 0; i < nd.actions.size(); i++) {
                double score = nd.Q[i] + nd.u[i];
                if (vl != 0.0) score = score - vl * (double)nd.virt[i];
                if (i == 0 || score > best_score) { best = (int)i; best_score = score; }
            }
            nd.virt[best] += 1;
            moves_out[j] = nd.actions[best];
        }
        t->path[j].push_back(Step{node, best});
        t->at[j] = child_of(t, node, best);
        alive++;
    }
    return alive;
}

int rag_tree_wave_backup(rag_tree* t, const double* values)
{
    if (!t || !values) return rag::fail(RAG_EINVAL, "null argument");
    for (int j = 0; j < t->count; j++) {
        for (const Step& s : t->path[j]) t->nodes[s.node].virt[s.i] -= 1;
        const double v = values[j];
        Node& root = t->nodes[t->root];
        root.visits += 1;
        root.q += (v - root.q) / (double)root.visits;
        for (const Step& s : t->path[j]) {
            Node& nd = t->nodes[s.node];
            nd.n[s.i] += 1;
            nd.Q[s.i] += (v - nd.Q[s.i]) / (double)nd.n[s.i];
            const long long parent_visits = visits_of(t, s.node);
            nd.u[s.i] = t->c_puct * nd.P[s.i] * sqrt((double)parent_visits) / (double)(1 + nd.n[s.i]);
        }
    }
    t->count = 0;
    return RAG_OK;
}

// update_with_move (mcts.py:208-217): the child for `action` becomes the root and keeps its subtree; anything
// else starts a new tree.  Returns 1 if a subtree was kept.
int rag_tree_advance(rag_tree* t, int action)
{
    if (!t) return rag::fail(RAG_EINVAL, "null tree");
    const Node& r = t->nodes[t->root];
    for (size_t i = 0; i < r.actions.size(); i++)
        if (r.actions[i] == action) {
            std::vector<Node> fresh;
            if (r.child[i] >= 0) copy_subtree(t->nodes, r.child[i], fresh, -1, 0);
            else fresh.push_back(Node());
            fresh[0].visits = r.n[i];
            fresh[0].q = r.Q[i];
            t->nodes.swap(fresh);
            t->root = 0; t->count = 0;
            return 1;
        }
    rag_tree_reset(t);
    return 0;
}

int rag_tree_root(const rag_tree* t) { return t ? t->root : -1; }
long long rag_tree_size(const rag_tree* t) { return t ? (long long)t->nodes.size() : 0; }

// node inspection: number of children (0 for a leaf), own counters, arrays (any pointer may be null)
int rag_tree_node_children(const rag_tree* t, int node)
{
    if (!t || node < 0 || node >= (int)t->nodes.size()) return rag::fail(RAG_EINVAL, "no such node");
    return (int)t->nodes[node].actions.size();
}

int rag_tree_node_stats(const rag_tree* t, int node, long long* visits, double* q, int32_t* parent, int32_t* index)
{
    if (!t || node < 0 || node >= (int)t->nodes.size()) return rag::fail(RAG_EINVAL, "no such node");
    const Node& nd = t->nodes[node];
    if (visits) *visits = visits_of(t, node);
    if (q) *q = nd.parent < 0 ? nd.q : t->nodes[nd.parent].Q[nd.index];
    if (parent) *parent = nd.parent;
    if (index) *index = nd.index;
    return RAG_OK;
}

int rag_tree_node_arrays(const rag_tree* t, int node, int32_t* actions, double* P, double* Q, double* u, int64_t* n,
                         int64_t* virt, int32_t* child)
{
    if (!t || node < 0 || node >= (int)t->nodes.size()) return rag::fail(RAG_EINVAL, "no such node");
    const Node& nd = t->nodes[node];
    for (size_t i = 0; i < nd.actions.size(); i++) {
        if (actions) actions[i] = nd.actions[i];
        if (P) P[i] = nd.P[i];
        if (Q) Q[i] = nd.Q[i];
        if (u) u[i] = nd.u[i];
        if (n) n[i] = nd.n[i];
        if (virt) virt[i] = nd.virt[i];
        if (child) child[i] = nd.child[i];
    }
    return RAG_OK;
}

}  // extern "C"
