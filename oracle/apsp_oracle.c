/*
 * ORACLE -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C restatement of the reference's Cython graph algorithms
 * (/root/reference/mDT/src/data/algos.pyx).  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load this library; the
 * product path (multimodaldiscussiontransformer_b200/) never does.
 *
 * Parity pin: checked bit-for-bit against the compiled reference algos.pyx
 * (oracle/_ref, built by oracle/build_ref.py from the file where it lies under
 * /root/reference) in tests/test_oracle_pinned.py and against the committed
 * fixtures tests/golden/apsp_*.npz that the same compiled reference produced.
 *
 * Build: make -C oracle   (gcc -O2 -shared -fPIC)
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define MDT_UNREACHABLE 510 /* algos.pyx:24-30,45-50 */

/*
 * floyd_warshall  (algos.pyx:7-52)
 *   adj  : [n,n] int64, row major (the reference casts any integer/bool matrix to long)
 *   M    : [n,n] int64 out -- min(shortest distance, 510)
 *   path : [n,n] int64 out -- last k (ascending) that strictly improved (i,j), -1 if none,
 *                             510 where M == 510
 */
void mdt_oracle_floyd_warshall(const int64_t *adj, int64_t *M, int64_t *path, int64_t n)
{
    int64_t i, j, k;
    memcpy(M, adj, (size_t)(n * n) * sizeof(int64_t));          /* :13-15 */
    for (i = 0; i < n * n; ++i) path[i] = -1;                   /* :16 */
    for (i = 0; i < n; ++i)                                     /* :25-30 */
        for (j = 0; j < n; ++j) {
            if (i == j) M[i * n + j] = 0;
            else if (M[i * n + j] == 0) M[i * n + j] = MDT_UNREACHABLE;
        }
    for (k = 0; k < n; ++k) {                                   /* :33-43 */
        const int64_t *Mk = M + n * k;
        for (i = 0; i < n; ++i) {
            int64_t *Mi = M + n * i;
            const int64_t Mik = Mi[k];
            for (j = 0; j < n; ++j) {
                const int64_t cost = Mik + Mk[j];
                if (Mi[j] > cost) {
                    Mi[j] = cost;
                    path[i * n + j] = k;
                }
            }
        }
    }
    for (i = 0; i < n * n; ++i)                                 /* :46-50 */
        if (M[i] >= MDT_UNREACHABLE) {
            path[i] = MDT_UNREACHABLE;
            M[i] = MDT_UNREACHABLE;
        }
}

/* get_all_edges (algos.pyx:55-60): in-order expansion of the predecessor matrix.
 * Returns the number of intermediate nodes written to out (caller sizes out >= n). */
static int64_t all_edges(const int64_t *path, int64_t n, int64_t i, int64_t j, int64_t *out, int64_t cap)
{
    const int64_t k = path[i * n + j];
    int64_t a, b;
    if (k == -1) return 0;
    a = all_edges(path, n, i, k, out, cap);
    if (a < cap) out[a] = k;
    b = all_edges(path, n, k, j, out + a + 1, cap - a - 1);
    return a + 1 + b;
}

/*
 * gen_edge_input  (algos.pyx:63-89)
 *   path      : [n,n] int64 from floyd_warshall
 *   edge_feat : [n,n,F] int64
 *   out       : [n,n,max_dist,F] int64, -1 filled; out[i,j,h,:] = edge_feat[p_h, p_{h+1}, :]
 * Returns 0, or -1 if a path is longer than max_dist (the reference would raise an
 * IndexError there).
 */
int mdt_oracle_gen_edge_input(int64_t max_dist, const int64_t *path, const int64_t *edge_feat,
                              int64_t *out, int64_t n, int64_t F)
{
    int64_t i, j, h, f;
    int64_t *nodes = (int64_t *)malloc((size_t)(n + 2) * sizeof(int64_t));
    int rc = 0;
    for (i = 0; i < n * n * max_dist * F; ++i) out[i] = -1;     /* :75 */
    for (i = 0; i < n; ++i)
        for (j = 0; j < n; ++j) {
            int64_t cnt, num_path;
            if (i == j) continue;                               /* :80-81 */
            if (path[i * n + j] == MDT_UNREACHABLE) continue;   /* :82-83 */
            nodes[0] = i;
            cnt = all_edges(path, n, i, j, nodes + 1, n);
            nodes[cnt + 1] = j;
            num_path = cnt + 1;                                 /* :85 */
            if (num_path > max_dist) { rc = -1; num_path = max_dist; }
            for (h = 0; h < num_path; ++h)                      /* :86-87 */
                for (f = 0; f < F; ++f)
                    out[((i * n + j) * max_dist + h) * F + f] =
                        edge_feat[(nodes[h] * n + nodes[h + 1]) * F + f];
        }
    free(nodes);
    return rc;
}

/* Batch driver used by the CPU-baseline timer: B matrices of the same n, back to back. */
void mdt_oracle_floyd_warshall_batch(const int64_t *adj, int64_t *M, int64_t *path, int64_t n, int64_t B)
{
    int64_t b;
    for (b = 0; b < B; ++b)
        mdt_oracle_floyd_warshall(adj + b * n * n, M + b * n * n, path + b * n * n, n);
}
