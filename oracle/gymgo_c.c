/* Oracle: independent C restatement of the GymGo rules (TEST INFRASTRUCTURE).
 *
 * PARITY UNPINNED: huangeddie/GymGo @ 3a3d97b is absent from /root/reference
 * (.SUBMODULES.json:9-15).  This file restates the same published algorithm as
 * oracle/gymgo.py (gym_go/gogame.py next_state/areas/winning,
 * gym_go/state_utils.py compute_invalid_moves/update_pieces/adj_data) but with a
 * different technique (explicit BFS group enumeration instead of
 * scipy.ndimage), so the two restatements can be differentially tested and
 * the CUDA kernels can be checked at full size (4096 boards x 722 plies).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
 * load the library built from this file (oracle/_build/libgymgo_c.so).
 *
 * State layout: uint8 planes [6][n][n]  (0 black, 1 white, 2 turn, 3 invalid,
 * 4 prev-pass, 5 done) -- the reference's planes (main.py:261) as bytes.
 */
#include <stdint.h>
#include <string.h>

#define MAXN 25
#define MAXP (MAXN * MAXN)

typedef struct {
    int n, np;
    int label[MAXP];   /* group id per point, -1 for none */
    int nlib[MAXP];    /* liberties per group id */
    int gsize[MAXP];
    int ngroups;
} Groups;

static int neighbours(int n, int p, int out[4]) {
    int r = p / n, c = p % n, k = 0;
    if (r > 0) out[k++] = p - n;
    if (r < n - 1) out[k++] = p + n;
    if (c > 0) out[k++] = p - 1;
    if (c < n - 1) out[k++] = p + 1;
    return k;
}

/* Enumerate groups of plane `stones`; liberties counted against `occupied`. */
static void find_groups(int n, const uint8_t* stones, const uint8_t* occupied, Groups* g) {
    int np = n * n, stack[MAXP], libmark[MAXP];
    g->n = n; g->np = np; g->ngroups = 0;
    for (int i = 0; i < np; i++) { g->label[i] = -1; libmark[i] = -1; }
    for (int s = 0; s < np; s++) {
        if (!stones[s] || g->label[s] >= 0) continue;
        int id = g->ngroups++, sp = 0, libs = 0, size = 0;
        stack[sp++] = s; g->label[s] = id;
        while (sp) {
            int p = stack[--sp], nb[4];
            size++;
            int k = neighbours(n, p, nb);
            for (int j = 0; j < k; j++) {
                int q = nb[j];
                if (stones[q]) {
                    if (g->label[q] < 0) { g->label[q] = id; stack[sp++] = q; }
                } else if (!occupied[q]) {
                    if (libmark[q] != id) { libmark[q] = id; libs++; }
                }
            }
        }
        g->nlib[id] = libs; g->gsize[id] = size;
    }
}

static void compute_invalid(uint8_t* st, int n, int player, int ko) {
    int np = n * n;
    uint8_t* own = st + player * np;
    uint8_t* opp = st + (1 - player) * np;
    uint8_t* inv = st + 3 * np;
    uint8_t occ[MAXP];
    Groups go, gp;
    for (int i = 0; i < np; i++) occ[i] = own[i] | opp[i];
    find_groups(n, own, occ, &go);
    find_groups(n, opp, occ, &gp);
    for (int p = 0; p < np; p++) {
        if (occ[p]) { inv[p] = 1; continue; }
        int nb[4], k = neighbours(n, p, nb);
        int surrounded = 1, possible = 0, definite = 0;
        for (int j = 0; j < k; j++) {
            int q = nb[j];
            if (!occ[q]) { surrounded = 0; continue; }
            if (own[q]) {
                int l = go.nlib[go.label[q]];
                if (l > 1) possible = 1; else if (l == 1) definite = 1;
            } else {
                int l = gp.nlib[gp.label[q]];
                if (l == 1) possible = 1; else if (l > 1) definite = 1;
            }
        }
        inv[p] = (uint8_t)(surrounded && possible && !definite);
    }
    if (ko >= 0) inv[ko] = 1;
}

/* returns 0 ok, -1 illegal point, -2 game already done */
int gymgo_c_next_state(uint8_t* st, int n, int action) {
    int np = n * n;
    if (st[5 * np]) return -2;
    int player = st[2 * np] ? 1 : 0;
    int prev_pass = st[4 * np];
    int ko = -1;
    if (action == np) {
        memset(st + 4 * np, 1, np);
        if (prev_pass) memset(st + 5 * np, 1, np);
    } else {
        if (action < 0 || action > np || st[3 * np + action]) return -1;
        memset(st + 4 * np, 0, np);
        uint8_t* own = st + player * np;
        uint8_t* opp = st + (1 - player) * np;
        own[action] = 1;
        int nb[4], k = neighbours(n, action, nb);
        int surrounded = 1;
        for (int j = 0; j < k; j++) if (!opp[nb[j]]) surrounded = 0;
        uint8_t occ[MAXP];
        for (int i = 0; i < np; i++) occ[i] = own[i] | opp[i];
        Groups g;
        find_groups(n, opp, occ, &g);
        int killed_ids[4], nk = 0, last_size = 0, last_point = -1;
        for (int j = 0; j < k; j++) {
            int q = nb[j];
            if (!opp[q]) continue;
            int id = g.label[q], seen = 0;
            for (int t = 0; t < nk; t++) if (killed_ids[t] == id) seen = 1;
            if (seen || g.nlib[id] > 0) continue;
            killed_ids[nk++] = id;
            last_size = g.gsize[id];
            last_point = q;
        }
        for (int t = 0; t < nk; t++)
            for (int p = 0; p < np; p++)
                if (g.label[p] == killed_ids[t]) opp[p] = 0;
        if (nk == 1 && surrounded && last_size == 1) ko = last_point;
    }
    compute_invalid(st, n, player, ko);
    memset(st + 2 * np, player ? 0 : 1, np);
    return 0;
}

void gymgo_c_areas(const uint8_t* st, int n, int* black_area, int* white_area) {
    int np = n * n, b = 0, w = 0, seen[MAXP], stack[MAXP];
    const uint8_t* B = st; const uint8_t* W = st + np;
    for (int i = 0; i < np; i++) { seen[i] = 0; b += B[i]; w += W[i]; }
    for (int s = 0; s < np; s++) {
        if (B[s] || W[s] || seen[s]) continue;
        int sp = 0, size = 0, tb = 0, tw = 0;
        stack[sp++] = s; seen[s] = 1;
        while (sp) {
            int p = stack[--sp], nb[4];
            size++;
            int k = neighbours(n, p, nb);
            for (int j = 0; j < k; j++) {
                int q = nb[j];
                if (B[q]) tb = 1; else if (W[q]) tw = 1;
                else if (!seen[q]) { seen[q] = 1; stack[sp++] = q; }
            }
        }
        if (tb && !tw) b += size; else if (tw && !tb) w += size;
    }
    *black_area = b; *white_area = w;
}

static uint64_t mix64(uint64_t x) {
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull;
    x ^= x >> 27; x *= 0x94d049bb133111ebull;
    x ^= x >> 31;
    return x;
}

/* Order-independent hash of a GymGo state, the same function as mzgo_board_hash in
 * muzero_go_b200/csrc/rules.cu: sum over rows of mix64(black | white<<20 | invalid<<40 | (row+1)<<59)
 * plus mix64(0xabcd + turn + 2*pass + 4*done).  Lets 4096x722-ply playouts be compared ply by ply. */
uint64_t gymgo_c_hash(const uint8_t* st, int n) {
    int np = n * n;
    uint64_t h = 0;
    for (int r = 0; r < n; r++) {
        uint64_t bl = 0, wh = 0, inv = 0;
        for (int c = 0; c < n; c++) {
            bl |= (uint64_t)(st[r * n + c] != 0) << c;
            wh |= (uint64_t)(st[np + r * n + c] != 0) << c;
            inv |= (uint64_t)(st[3 * np + r * n + c] != 0) << c;
        }
        h += mix64(bl | (wh << 20) | (inv << 40) | ((uint64_t)(r + 1) << 59));
    }
    h += mix64(0xabcdull + (uint64_t)(st[2 * np] != 0) + 2ull * (st[4 * np] != 0) + 4ull * (st[5 * np] != 0));
    return h;
}

/* Config-3 playout: for each board, ply t plays the k-th (ascending) entry of
 * {valid board points} U {pass}, k = R[b*T + t] % n_valid.  Stops a board when
 * done.  hashes (optional) [nboards][T] gymgo_c_hash after every ply.  trace (optional) receives the state after every ply:
 * [nboards][T][6*n*n] bytes; actions (optional) [nboards][T] int32 (-1 = not played).
 * Returns total plies played. */
long gymgo_c_playout(uint8_t* states, int n, int nboards, const uint32_t* R, int T,
                     uint8_t* trace, int32_t* actions, uint64_t* hashes) {
    int np = n * n, S = 6 * np;
    long plies = 0;
    for (int b = 0; b < nboards; b++) {
        uint8_t* st = states + (long)b * S;
        for (int t = 0; t < T; t++) {
            int a = -1;
            if (!st[5 * np]) {
                int nvalid = 1;
                for (int p = 0; p < np; p++) nvalid += !st[3 * np + p];
                int k = (int)(R[(long)b * T + t] % (uint32_t)nvalid);
                a = np;
                for (int p = 0; p < np; p++)
                    if (!st[3 * np + p]) { if (k == 0) { a = p; break; } k--; }
                gymgo_c_next_state(st, n, a);
                plies++;
            }
            if (actions) actions[(long)b * T + t] = a;
            if (trace) memcpy(trace + ((long)b * T + t) * S, st, S);
            if (hashes) hashes[(long)b * T + t] = gymgo_c_hash(st, n);
        }
    }
    return plies;
}
