/* Oracle: a THIRD rules checker, written from the rules of Go themselves (TEST INFRASTRUCTURE).
 *
 * oracle/gymgo.py and oracle/gymgo_c.c both restate GymGo's liberty-count FORMULA for the invalid-move plane
 * (possibly_invalid / definitely_valid / surrounded) -- one recollection written twice.  This file does not know
 * that formula.  It plays Go the way a rule book states it:
 *
 *   1. a move places a stone of the mover's colour on an empty point;
 *   2. then every group of the OPPONENT that has no liberty left is removed;
 *   3. the move is illegal if the mover's own group then has no liberty (suicide);
 *   4. simple ko: if the move captured exactly one stone and the capturing stone is then a lone stone with exactly
 *      one liberty, the opponent may not play on the captured point on the very next ply (a pass lifts the ban too);
 *   5. two consecutive passes end the game; area scoring: a colour's area = its stones + the empty regions that
 *      touch only that colour.
 *
 * From that it derives the six GymGo planes the reference reads (main.py:261, :637, :681; govars order black,
 * white, turn, invalid, pass, done): invalid = occupied | illegal-by-play-and-check | ko point, for the player to
 * move.  Agreement of these planes with oracle/gymgo_c.c on every position of a playout says: the restated
 * formula IS textbook legality + GymGo's plane conventions.  (GymGo itself -- huangeddie/GymGo@3a3d97b -- is absent
 * from /root/reference and from the GPU box, profiles/r02/gymgo_probe.txt: rules parity stays "unpinned by the
 * reference".)
 *
 * Only tests/ may load the library built from this file (oracle/_build/libgo_textbook.so).
 */
#include <stdint.h>
#include <string.h>

#define MAXN 25
#define MAXP (MAXN * MAXN)

typedef struct {
    int n;
    int8_t pt[MAXP]; /* 0 empty, 1 black, 2 white */
    int to_move;     /* 0 black, 1 white */
    int ko;          /* point banned for the player to move, -1 none */
    int last_pass;   /* previous ply was a pass */
    int over;
} Game;

static int nbrs(int n, int p, int* out) {
    int r = p / n, c = p - r * n, k = 0;
    if (r > 0) out[k++] = p - n;
    if (c > 0) out[k++] = p - 1;
    if (c < n - 1) out[k++] = p + 1;
    if (r < n - 1) out[k++] = p + n;
    return k;
}

/* The group of the stone at `start`: members into grp[], returns its size; *libs = number of distinct liberties. */
static int group_of(const int8_t* pt, int n, int start, int* grp, int* libs) {
    uint8_t seen[MAXP];
    memset(seen, 0, (size_t)n * n);
    int8_t col = pt[start];
    int size = 0, head = 0, nl = 0;
    grp[size++] = start;
    seen[start] = 1;
    while (head < size) {
        int p = grp[head++], nb[4], k = nbrs(n, p, nb);
        for (int j = 0; j < k; j++) {
            int q = nb[j];
            if (seen[q]) continue;
            if (pt[q] == col) { seen[q] = 1; grp[size++] = q; }
            else if (pt[q] == 0) { seen[q] = 1; nl++; }
        }
    }
    *libs = nl;
    return size;
}

/* Play colour `col` (1/2) at empty point p on board pt (modified).  Returns -1 if the move is suicide (board then
 * undefined), else the number of stones captured; *cap_pt = one captured point; *own_size / *own_libs describe the
 * mover's group after the captures. */
static int place(int8_t* pt, int n, int p, int8_t col, int* cap_pt, int* own_size, int* own_libs) {
    int grp[MAXP], libs, nb[4], captured = 0;
    pt[p] = col;
    int k = nbrs(n, p, nb);
    for (int j = 0; j < k; j++) {
        int q = nb[j];
        if (pt[q] != 3 - col) continue;           /* (a group removed through an earlier neighbour is empty by now) */
        int size = group_of(pt, n, q, grp, &libs);
        if (libs == 0) {
            for (int i = 0; i < size; i++) pt[grp[i]] = 0;
            captured += size;
            *cap_pt = q;
        }
    }
    *own_size = group_of(pt, n, p, grp, own_libs);
    return *own_libs == 0 ? -1 : captured;
}

static int legal(const Game* g, int p) {
    if (g->pt[p] != 0 || p == g->ko) return 0;
    int nb[4], k = nbrs(g->n, p, nb);
    /* an empty neighbour stays empty (captures only remove stones): the new stone has a liberty, no need to play it out */
    for (int j = 0; j < k; j++) if (g->pt[nb[j]] == 0) return 1;
    int8_t tmp[MAXP];
    int cap_pt, os, ol;
    memcpy(tmp, g->pt, (size_t)g->n * g->n);
    return place(tmp, g->n, p, (int8_t)(g->to_move + 1), &cap_pt, &os, &ol) >= 0;
}

/* literal play-and-check on every empty point, no shortcut (cross-checks `legal`) */
static int legal_literal(const Game* g, int p) {
    if (g->pt[p] != 0 || p == g->ko) return 0;
    int8_t tmp[MAXP];
    int cap_pt, os, ol;
    memcpy(tmp, g->pt, (size_t)g->n * g->n);
    return place(tmp, g->n, p, (int8_t)(g->to_move + 1), &cap_pt, &os, &ol) >= 0;
}

/* returns 0 ok, -1 illegal move, -2 game over */
static int play(Game* g, int action) {
    int np = g->n * g->n;
    if (g->over) return -2;
    if (action == np) {
        if (g->last_pass) g->over = 1;
        g->last_pass = 1;
        g->ko = -1;
    } else {
        if (action < 0 || action > np || !legal(g, action)) return -1;
        int cap_pt = -1, os, ol;
        int captured = place(g->pt, g->n, action, (int8_t)(g->to_move + 1), &cap_pt, &os, &ol);
        g->ko = (captured == 1 && os == 1 && ol == 1) ? cap_pt : -1;
        g->last_pass = 0;
    }
    g->to_move ^= 1;
    return 0;
}

static void planes_of(const Game* g, uint8_t* st, int literal) {
    int n = g->n, np = n * n;
    for (int p = 0; p < np; p++) {
        st[p] = g->pt[p] == 1;
        st[np + p] = g->pt[p] == 2;
        st[2 * np + p] = (uint8_t)g->to_move;
        st[3 * np + p] = (uint8_t)!(literal ? legal_literal(g, p) : legal(g, p));
        st[4 * np + p] = (uint8_t)g->last_pass;
        st[5 * np + p] = (uint8_t)g->over;
    }
}

static void game_from_planes(Game* g, const uint8_t* st, int n, int ko) {
    int np = n * n;
    g->n = n;
    for (int p = 0; p < np; p++) g->pt[p] = st[p] ? 1 : (st[np + p] ? 2 : 0);
    g->to_move = st[2 * np] ? 1 : 0;
    g->ko = ko;
    g->last_pass = st[4 * np] != 0;
    g->over = st[5 * np] != 0;
}

static void area_score(const Game* g, int* black, int* white) {
    int n = g->n, np = n * n, b = 0, w = 0, queue[MAXP];
    uint8_t seen[MAXP];
    memset(seen, 0, (size_t)np);
    for (int s = 0; s < np; s++) {
        if (g->pt[s] == 1) b++;
        else if (g->pt[s] == 2) w++;
        else if (!seen[s]) {
            int head = 0, size = 0, touch = 0;
            queue[size++] = s; seen[s] = 1;
            while (head < size) {
                int p = queue[head++], nb[4], k = nbrs(n, p, nb);
                for (int j = 0; j < k; j++) {
                    int q = nb[j];
                    if (g->pt[q]) touch |= g->pt[q];
                    else if (!seen[q]) { seen[q] = 1; queue[size++] = q; }
                }
            }
            if (touch == 1) b += size;
            else if (touch == 2) w += size;
        }
    }
    *black = b; *white = w;
}

/* The state hash the CUDA kernels and oracle/gymgo_c.c print per ply (mzgo_board_hash in csrc/rules.cu): a
 * checksum of the six planes, no rules in it. */
static uint64_t mix64(uint64_t x) {
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull;
    x ^= x >> 27; x *= 0x94d049bb133111ebull;
    x ^= x >> 31;
    return x;
}
static uint64_t planes_hash(const uint8_t* st, int n) {
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

/* ------------------------------------------------------------------------------------------------ exported */

/* One move on a position given as planes.  `ko_in` = the point banned for the mover (-1 none; the planes alone do not
 * say which invalid point is the ko point).  Writes the planes after the move, returns 0 / -1 / -2; *ko_out = the
 * new ban. */
int tb_next_state(const uint8_t* st_in, int n, int ko_in, int action, uint8_t* st_out, int* ko_out, int literal) {
    Game g;
    game_from_planes(&g, st_in, n, ko_in);
    int rc = play(&g, action);
    if (rc) return rc;
    planes_of(&g, st_out, literal);
    *ko_out = g.ko;
    return 0;
}

void tb_areas(const uint8_t* st, int n, int* black, int* white) {
    Game g;
    game_from_planes(&g, st, n, -1);
    area_score(&g, black, white);
}

/* BASELINE.json configs[2]: ply t of board b plays the k-th (ascending) entry of {legal points} U {pass},
 * k = R[b*T+t] % count; finished boards stop.  Everything from this engine's own rules.  Outputs (optional):
 * actions [B][T] (-1 = not played), hashes [B][T] of the planes after every ply, final planes [B][6][n][n],
 * areas [B][2] of the final position.  Returns the number of plies played. */
long tb_playout(int n, int nboards, const uint32_t* R, int T, int32_t* actions, uint64_t* hashes, uint8_t* final_planes,
                int32_t* areas, int literal) {
    int np = n * n;
    long plies = 0;
    uint8_t st[6 * MAXP];
    for (int b = 0; b < nboards; b++) {
        Game g;
        memset(&g, 0, sizeof(g));
        g.n = n; g.ko = -1;
        planes_of(&g, st, literal);
        for (int t = 0; t < T; t++) {
            int a = -1;
            if (!g.over) {
                int count = 1;
                for (int p = 0; p < np; p++) count += !st[3 * np + p];
                int k = (int)(R[(long)b * T + t] % (uint32_t)count);
                a = np;
                for (int p = 0; p < np; p++)
                    if (!st[3 * np + p]) { if (k == 0) { a = p; break; } k--; }
                play(&g, a);
                planes_of(&g, st, literal);
                plies++;
            }
            if (actions) actions[(long)b * T + t] = a;
            if (hashes) hashes[(long)b * T + t] = planes_hash(st, n);
        }
        if (final_planes) memcpy(final_planes + (long)b * 6 * np, st, (size_t)6 * np);
        if (areas) area_score(&g, &areas[2 * b], &areas[2 * b + 1]);
    }
    return plies;
}
