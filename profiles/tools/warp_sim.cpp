// warp_sim.cpp -- CPU model of how a warp of the extend kernel spends its issue slots (profiles/tools/warp_sim.py drives it).
//
// Not product code and not the oracle: a planning tool.  It walks the SAME binary tree arrays the kernels traverse (CudaBVH
// compact layout, CudaBVH.cpp:196-224) with the same per-ray state machine (one inner-node visit / one triangle test per step,
// first leaf postponed, near child first: cudaUtils.h:298-448), for real secondary rays of the scene (diffuse bounces off the
// mesh and the sphere walls), and replays 32 rays at a time under a chosen warp SCHEDULE.  Output: warp-level steps per ray and
// lanes active per step -- the two numbers ncu reports for the real kernel (smsp__inst_executed per ray,
// smsp__thread_inst_executed_per_inst_executed) -- so that schedules can be compared before any CUDA is written.
//
//   warp_sim <scene.bin> <width> <height> <bounces> <policy> [node_k] [refill_min] [rays_per_lane]
// scene.bin: int32 n_node_vec4, n_tri_vec4, n_spheres; float nodes[4*n_node_vec4], tris[4*n_tri_vec4], spheres[11*n_spheres]
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <queue>
#include <vector>

struct V3 { float x, y, z; };
static inline V3 v3(float x, float y, float z) { return {x, y, z}; }
static inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
static inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
static inline V3 norm(V3 a) { float l = std::sqrt(dot(a, a)); return a * (1.0f / l); }

static std::vector<float> g_nodes, g_tris, g_sph;
static int g_n_sph = 0;
static const int SENTINEL = 0x76543210;

struct Ray { V3 o, d; };
struct Trav {      // per-lane state, as TravState in gpt_traverse.cuh (children are byte offsets / ~slot like the reference arrays)
    float idx, idy, idz, oodx, oody, oodz, t;
    int tri, node, leaf, tri_cur, sp;
    int stack[64];
    bool active;
    int ray;
    int leafq[8], nleaf;      // "m"/"p" schedules: up to M postponed leaves per lane (slot of the first triangle of each)
};
static inline bool wants_node(const Trav& s) { return s.active && (unsigned)s.node < (unsigned)SENTINEL && s.node >= 0; }
static inline bool wants_tri(const Trav& s) { return s.active && s.leaf < 0; }
static inline bool finished(const Trav& s) { return s.node == SENTINEL && s.leaf >= 0; }

static void trav_init(Trav& s, const Ray& r, int ray)
{
    auto guard = [](float d) { const float e = 8.2718061255302767e-25f; return std::fabs(d) > e ? d : std::copysign(e, d); };
    s.idx = 1.0f / guard(r.d.x); s.idy = 1.0f / guard(r.d.y); s.idz = 1.0f / guard(r.d.z);
    s.oodx = r.o.x * s.idx; s.oody = r.o.y * s.idy; s.oodz = r.o.z * s.idz;
    s.t = 3.402823466e+38f; s.tri = -1; s.sp = 0; s.stack[0] = SENTINEL; s.node = 0; s.leaf = 0; s.tri_cur = 0; s.active = true; s.ray = ray; s.nleaf = 0;
}
static void node_step(Trav& s)
{
    const float* n = &g_nodes[(size_t)(s.node / 16) * 4];      // byte offset -> float index
    const float c0lox = n[0] * s.idx - s.oodx, c0hix = n[1] * s.idx - s.oodx, c0loy = n[2] * s.idy - s.oody, c0hiy = n[3] * s.idy - s.oody;
    const float c1lox = n[4] * s.idx - s.oodx, c1hix = n[5] * s.idx - s.oodx, c1loy = n[6] * s.idy - s.oody, c1hiy = n[7] * s.idy - s.oody;
    const float c0loz = n[8] * s.idz - s.oodz, c0hiz = n[9] * s.idz - s.oodz, c1loz = n[10] * s.idz - s.oodz, c1hiz = n[11] * s.idz - s.oodz;
    int c0, c1; memcpy(&c0, n + 12, 4); memcpy(&c1, n + 13, 4);
    const float c0min = std::max(std::max(std::min(c0lox, c0hix), std::min(c0loy, c0hiy)), std::max(std::min(c0loz, c0hiz), 0.0f));
    const float c0max = std::min(std::min(std::max(c0lox, c0hix), std::max(c0loy, c0hiy)), std::min(std::max(c0loz, c0hiz), s.t));
    const float c1min = std::max(std::max(std::min(c1lox, c1hix), std::min(c1loy, c1hiy)), std::max(std::min(c1loz, c1hiz), 0.0f));
    const float c1max = std::min(std::min(std::max(c1lox, c1hix), std::max(c1loy, c1hiy)), std::min(std::max(c1loz, c1hiz), s.t));
    const bool t0 = c0min <= c0max, t1 = c1min <= c1max;
    int next;
    if (!t0 && !t1) { next = s.stack[s.sp]; --s.sp; }
    else {
        next = t0 ? c0 : c1;
        if (t0 && t1) { int far_c = c1; if (c1min < c0min) { far_c = next; next = c1; } s.stack[++s.sp] = far_c; }
    }
    if (next < 0 && s.leaf >= 0) { s.leaf = next; s.tri_cur = ~next; next = s.stack[s.sp]; --s.sp; }
    s.node = next;
}
// one triangle of the postponed leaf; returns true when the leaf is finished
static bool tri_step(Trav& s, const Ray& r)
{
    const float* v = &g_tris[(size_t)s.tri_cur * 4];
    uint32_t bits; memcpy(&bits, v, 4);
    if (bits == 0x80000000u) {              // terminator: counted as part of the previous step by the callers (we never step onto it)
        s.leaf = s.node;
        if (s.node < 0) { s.tri_cur = ~s.node; s.node = s.stack[s.sp]; --s.sp; }
        return true;
    }
    const V3 v0 = v3(v[0], v[1], v[2]), v1 = v3(v[4], v[5], v[6]), v2 = v3(v[8], v[9], v[10]);
    const V3 e1 = v1 - v0, e2 = v2 - v0, pv = cross(r.d, e2);
    const float det = dot(e1, pv);
    if (det >= 0.00001f) {                  // culling on
        const float inv = 1.0f / det; const V3 tv = r.o - v0; const float u = dot(tv, pv) * inv;
        if (u >= 0.0f && u <= 1.0f) {
            const V3 qv = cross(tv, e1); const float vv = dot(r.d, qv) * inv;
            if (vv >= 0.0f && u + vv <= 1.0f) { const float t = dot(e2, qv) * inv; if (t > 0.0f && t < s.t) { s.t = t; s.tri = s.tri_cur; } }
        }
    }
    s.tri_cur += 3;
    memcpy(&bits, &g_tris[(size_t)s.tri_cur * 4], 4);
    if (bits == 0x80000000u) {
        s.leaf = s.node;
        if (s.node < 0) { s.tri_cur = ~s.node; s.node = s.stack[s.sp]; --s.sp; }
        return true;
    }
    return false;
}
// ---- multi-leaf variant: a lane keeps descending while it has room for another postponed leaf
static inline bool m_wants_node(const Trav& s) { return s.active && s.node >= 0 && s.node != SENTINEL; }
static inline bool m_finished(const Trav& s) { return s.node == SENTINEL && s.nleaf == 0; }
static void node_step_m(Trav& s, int M)
{
    s.leaf = s.nleaf < M ? 0 : -1;       // room left: node_step postpones the leaf it meets (moved to the queue below); else it stalls on it
    node_step(s);
    if (s.nleaf < M && s.leaf < 0) s.leafq[s.nleaf++] = ~s.leaf;
    s.leaf = 0;
    while (s.node < 0 && s.nleaf < M) { s.leafq[s.nleaf++] = ~s.node; s.node = s.stack[s.sp]; --s.sp; }
    // node < 0 with a full queue: stalled until the triangle phase
}
static int leaf_size(int slot) { int n = 0; uint32_t b; for (;; ++n) { memcpy(&b, &g_tris[(size_t)(slot + 3 * n) * 4], 4); if (b == 0x80000000u) break; } return n; }
static void test_leaf(Trav& s, const Ray& r, int slot)
{
    Trav tmp = s; tmp.leaf = -1; tmp.tri_cur = slot; tmp.node = SENTINEL;      // reuse tri_step's arithmetic on a scratch copy
    uint32_t b; memcpy(&b, &g_tris[(size_t)slot * 4], 4);
    if (b != 0x80000000u) while (!tri_step(tmp, r)) {}
    s.t = tmp.t; s.tri = tmp.tri;
}
static void after_tri_phase(Trav& s)
{
    s.nleaf = 0;
    while (s.node < 0) { s.leafq[s.nleaf++] = ~s.node; s.node = s.stack[s.sp]; --s.sp; if (s.nleaf >= 1) break; }
}

static void trace_full(const Ray& r, float& t, int& tri)
{
    Trav s; trav_init(s, r, 0);
    while (!finished(s)) { while (wants_node(s)) node_step(s); while (wants_tri(s)) tri_step(s, r); }
    t = s.t; tri = s.tri;
}

// ---- path generator: secondary rays of a diffuse scene (mesh DIFF, spheres DIFF), reference lobe (cudaUtils.h:185-192)
static uint64_t g_rng = 0x9E3779B97F4A7C15ull;
static inline float rnd() { g_rng ^= g_rng << 13; g_rng ^= g_rng >> 7; g_rng ^= g_rng << 17; return (float)((g_rng >> 40) + 1) * (1.0f / 16777216.0f); }
static bool bounce(Ray& r)
{
    float t; int tri; trace_full(r, t, tri);
    int geom = tri >= 0 ? 0 : 3, prim = tri;
    for (int i = 0; i < g_n_sph; ++i) {
        const float* s = &g_sph[11 * i];
        const V3 op = v3(s[0], s[1], s[2]) - r.o; const float b = dot(op, r.d); float disc = b * b - dot(op, op) + s[3] * s[3];
        if (disc < 0) continue;
        disc = std::sqrt(disc);
        const float ts = b - disc > 0.01f ? b - disc : (b + disc > 0.01f ? b + disc : 0.0f);
        if (ts != 0.0f && ts < t && ts > 0.01f) { t = ts; geom = 1; prim = i; }
    }
    if (geom == 3) return false;
    const V3 hp = r.o + r.d * t;
    V3 nl;
    if (geom == 1) { const float* s = &g_sph[11 * prim]; V3 n = norm(hp - v3(s[0], s[1], s[2])); nl = dot(n, r.d) < 0 ? n : n * -1.0f; }
    else { const float* v = &g_tris[(size_t)prim * 4]; nl = norm(cross(v3(v[0], v[1], v[2]) - v3(v[4], v[5], v[6]), v3(v[0], v[1], v[2]) - v3(v[8], v[9], v[10]))); }
    V3 nt = std::fabs(nl.x) > std::fabs(nl.y) ? v3(nl.z, 0, -nl.x) : v3(0, -nl.z, nl.y); nt = norm(nt);
    const V3 nb = norm(cross(nl, nt));
    const float f1 = rnd(), f2 = rnd(), ph = 6.2831853f * f1;
    const V3 rv = v3(std::cos(ph) * f2, std::sqrt(std::max(0.0f, 1 - f2 * f2)), std::sin(ph) * f2);
    r.d = norm(nb * rv.x + nl * rv.y + nt * rv.z);
    r.o = hp + nl * 0.001f;
    return true;
}

// ---- warp schedules.  Costs in warp instructions, from the SASS of the round-1 kernel (profiles/r01_schedules.md):
static int C_NODE = 75, C_TRI = 70, C_REFILL = 45, C_ROUND = 12, C_VOTE = 4;
struct Stats { double slots = 0, lane_slots = 0, node_slots = 0, node_lanes = 0, tri_slots = 0, tri_lanes = 0, other = 0; long rays = 0; };

struct Warp {
    Trav l[32];
    double clock = 0;
    bool exhausted = false;
};

int main(int argc, char** argv)
{
    if (argc < 6) { fprintf(stderr, "usage: warp_sim scene.bin W H bounces policy [node_k refill_min rays_per_lane]\n"); return 2; }
    FILE* f = fopen(argv[1], "rb");
    if (!f) { perror("scene"); return 1; }
    int hdr[3];
    if (fread(hdr, 4, 3, f) != 3) return 1;
    g_nodes.resize((size_t)hdr[0] * 4); g_tris.resize((size_t)hdr[1] * 4 + 8); g_sph.resize((size_t)hdr[2] * 11); g_n_sph = hdr[2];
    if (fread(g_nodes.data(), 4, g_nodes.size(), f) != g_nodes.size()) return 1;
    if (fread(g_tris.data(), 4, (size_t)hdr[1] * 4, f) != (size_t)hdr[1] * 4) return 1;
    if (g_n_sph && fread(g_sph.data(), 4, g_sph.size(), f) != g_sph.size()) return 1;
    fclose(f);
    const int W = atoi(argv[2]), H = atoi(argv[3]), n_bounce = atoi(argv[4]);
    const char* policy = argv[5];
    const int node_k = argc > 6 ? atoi(argv[6]) : 3, refill_min = argc > 7 ? atoi(argv[7]) : 8;
    const int rays_per_lane = argc > 8 ? atoi(argv[8]) : 175;
    if (getenv("C_NODE")) C_NODE = atoi(getenv("C_NODE"));
    if (getenv("C_TRI")) C_TRI = atoi(getenv("C_TRI"));

    // camera: BasicScene defaults (pos 0, looking down -z, dist = H/60, fov = tan 45 = 1)
    std::vector<Ray> rays;
    const float dist = (float)(H / 60 > 0 ? H / 60 : 1), aspect = (float)W / H;
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            const float xs = ((float)x - W / 2.0f + rnd()) * dist * aspect / (W - 1), ys = ((float)y - H / 2.0f + rnd()) * dist / (H - 1);
            const V3 dir = v3(xs, ys, -dist);
            rays.push_back({dir, norm(dir)});
        }
    for (int b = 0; b < n_bounce; ++b) {
        // the bounce's queue = surviving paths in path order (the shade kernel's compaction keeps it nearly so)
        const size_t n = rays.size();
        // ---- simulate: n_warps persistent warps share the queue; a warp fetches when its clock is the smallest (event order)
        size_t n_warps = std::max<size_t>(1, n / ((size_t)32 * rays_per_lane));
        std::vector<Warp> warps(n_warps);
        for (auto& w : warps) for (auto& s : w.l) { s.active = false; s.node = SENTINEL; s.leaf = 0; }
        size_t cursor = 0;
        Stats st;
        auto cmp = [&](size_t a, size_t b2) { return warps[a].clock > warps[b2].clock; };
        std::priority_queue<size_t, std::vector<size_t>, decltype(cmp)> pq(cmp);
        for (size_t i = 0; i < n_warps; ++i) pq.push(i);
        double t_end = 0;
        auto spend = [&](Warp& w, int cost, int lanes, int kind) {
            w.clock += cost; st.slots += cost; st.lane_slots += (double)cost * lanes;
            if (kind == 0) { st.node_slots += cost; st.node_lanes += (double)cost * lanes; }
            else if (kind == 1) { st.tri_slots += cost; st.tri_lanes += (double)cost * lanes; }
            else st.other += cost;
        };
        auto refill = [&](Warp& w) {
            int idle = 0; for (auto& s : w.l) idle += !s.active;
            if (!idle || w.exhausted) return;
            int got = 0;
            for (auto& s : w.l) if (!s.active) { if (cursor < n) { trav_init(s, rays[cursor], (int)cursor); ++cursor; ++got; } }
            if (cursor >= n) w.exhausted = true;
            spend(w, C_REFILL, got, 2);
        };
        const bool in_loop = strstr(policy, "inloop") != nullptr;      // refill inside the node loop as soon as refill_min lanes idle
        const bool dual = strstr(policy, "ifif") != nullptr;           // per-step vote between node and tri work (upper bound)
        const bool tri_exit = strstr(policy, "triexit") != nullptr;    // leave the triangle loop when few lanes remain
        // "m<M>": up to M postponed leaves per lane, every lane tests its own; "p<M>": the same, triangle phase pooled over the warp.
        // node_k is then the number of pending LEAVES in the warp that starts a triangle phase
        int multi = 0; bool pooled = false;
        if ((policy[0] == 'm' || policy[0] == 'p') && policy[1] >= '1' && policy[1] <= '8') { multi = policy[1] - '0'; pooled = policy[0] == 'p'; }
        static int C_POOL = getenv("C_POOL") ? atoi(getenv("C_POOL")) : 20;         // list + ray fetch from shared memory, per pass
        static int C_PHASE = getenv("C_PHASE") ? atoi(getenv("C_PHASE")) : 90;      // scan + list build + result read-back, per phase
        static int C_MVOTE = getenv("C_MVOTE") ? atoi(getenv("C_MVOTE")) : 10;      // three ballots per node-loop trip
        static int HOLDERS = getenv("HOLDERS") ? atoi(getenv("HOLDERS")) : 0;       // > 0: trigger on lanes holding a leaf instead of pending leaves
        while (!pq.empty()) {
            const size_t wi = pq.top(); pq.pop();
            Warp& w = warps[wi];
            // ---- one round of the warp
            int idle = 0; for (auto& s : w.l) idle += !s.active;
            if (idle && !w.exhausted && (idle >= refill_min || idle == 32)) refill(w);
            int act = 0; for (auto& s : w.l) act += s.active;
            if (!act) { t_end = std::max(t_end, w.clock); continue; }      // warp done
            spend(w, C_ROUND, act, 2);
            if (multi) {
                while (true) {
                    int in = 0, pending = 0, stalled = 0, holders = 0;
                    for (auto& l : w.l) { in += m_wants_node(l); pending += l.active ? l.nleaf + (l.node < 0) : 0; stalled += l.active && l.node < 0; holders += l.active && (l.nleaf > 0 || l.node < 0); }
                    if (!in || (HOLDERS ? holders >= HOLDERS : pending >= node_k)) break;
                    for (auto& l : w.l) if (m_wants_node(l)) node_step_m(l, multi);
                    spend(w, C_NODE + C_MVOTE, in, 0);
                    for (auto& l : w.l) if (l.active && m_finished(l)) { l.active = false; ++st.rays; }
                    int id2 = 0; for (auto& l : w.l) id2 += !l.active;
                    if (id2 >= refill_min && !w.exhausted) refill(w);
                }
                // triangle phase
                int total = 0, mx = 0, holders = 0;
                int per[32];
                for (int i = 0; i < 32; ++i) {
                    per[i] = 0;
                    Trav& l = w.l[i];
                    if (!l.active) continue;
                    for (int q = 0; q < l.nleaf; ++q) { per[i] += leaf_size(l.leafq[q]); test_leaf(l, rays[l.ray], l.leafq[q]); }
                    total += per[i]; mx = std::max(mx, per[i]); holders += per[i] > 0;
                    after_tri_phase(l);
                }
                if (pooled) { for (int done = 0; done < total; done += 32) spend(w, C_TRI + C_POOL, std::min(32, total - done), 1); if (total) spend(w, C_PHASE, 32, 2); }
                else for (int it = 0; it < mx; ++it) { int ln = 0; for (int i = 0; i < 32; ++i) ln += per[i] > it; spend(w, C_TRI, ln, 1); }
                for (auto& l : w.l) if (l.active && m_finished(l)) { l.active = false; ++st.rays; }
                pq.push(wi);
                continue;
            }
            if (dual) {
                int nn = 0, nt = 0; for (auto& s : w.l) { nn += wants_node(s); nt += wants_tri(s); }
                if (nt >= nn && nt > 0) { for (auto& s : w.l) if (wants_tri(s)) tri_step(s, rays[s.ray]); spend(w, C_TRI + C_VOTE, nt, 1); }
                else if (nn > 0) { for (auto& s : w.l) if (wants_node(s)) node_step(s); spend(w, C_NODE + C_VOTE, nn, 0); }
            } else {
                // node loop
                while (true) {
                    int in = 0, lacking = 0;
                    for (auto& s : w.l) if (wants_node(s)) { ++in; }
                    if (!in) break;
                    for (auto& s : w.l) if (wants_node(s)) node_step(s);
                    spend(w, C_NODE + C_VOTE, in, 0);
                    for (auto& s : w.l) if (wants_node(s) && s.leaf >= 0) ++lacking;
                    if (strstr(policy, "hold")) {            // exit when node_k lanes hold a leaf (not: when few still lack one)
                        int holders = 0; for (auto& s : w.l) holders += s.active && s.leaf < 0;
                        if (holders >= node_k) break;
                    } else
                    if (lacking < node_k) break;
                    if (in_loop) {
                        for (auto& s : w.l) if (s.active && finished(s)) { s.active = false; ++st.rays; }
                        int id2 = 0; for (auto& s : w.l) id2 += !s.active;
                        if (id2 >= refill_min && !w.exhausted) refill(w);
                    }
                }
                // triangle loop: one leaf per lane per round
                bool busy[32]; int nb = 0;
                for (int i = 0; i < 32; ++i) { busy[i] = wants_tri(w.l[i]); nb += busy[i]; }
                while (nb) {
                    int stepped = nb;
                    for (int i = 0; i < 32; ++i) if (busy[i]) { if (tri_step(w.l[i], rays[w.l[i].ray])) { busy[i] = false; --nb; } }
                    spend(w, C_TRI, stepped, 1);
                    if (tri_exit && nb && nb < 4) break;
                }
            }
            for (auto& s : w.l) if (s.active && finished(s)) { s.active = false; ++st.rays; }
            pq.push(wi);
        }
        const double kernel_slots = t_end * (double)n_warps;      // issue slots the launch holds the SMs for (tail included)
        printf("{\"bounce\": %d, \"rays\": %zu, \"policy\": \"%s\", \"node_k\": %d, \"refill_min\": %d, \"warps\": %zu, "
               "\"warp_inst_per_ray\": %.1f, \"with_tail\": %.1f, \"lanes\": %.2f, \"node_share\": %.3f, \"node_lanes\": %.2f, \"tri_share\": %.3f, \"tri_lanes\": %.2f, \"other_share\": %.3f}\n",
               b, n, policy, node_k, refill_min, n_warps, st.slots / n, kernel_slots / n, st.lane_slots / st.slots, st.node_slots / st.slots,
               st.node_lanes / std::max(1.0, st.node_slots), st.tri_slots / st.slots, st.tri_lanes / std::max(1.0, st.tri_slots), st.other / st.slots);
        fflush(stdout);
        // ---- next bounce's rays
        std::vector<Ray> next; next.reserve(n);
        for (auto& r : rays) { Ray q = r; if (bounce(q)) next.push_back(q); }
        rays.swap(next);
        if (rays.empty()) break;
    }
    return 0;
}
