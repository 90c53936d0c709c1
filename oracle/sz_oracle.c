/*
 * sz_oracle.c -- TEST INFRASTRUCTURE ONLY (see sz_oracle.h): CPU restatement of the SZ 2.1.12
 * blocked-regression path.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
 * load the library built from this file.  The product (sz_b200/) never links or calls it.
 *
 * Build: gcc -O2 -fPIC -std=gnu99 -ffp-contract=off -shared sz_oracle.c -lm   (oracle/Makefile, target "port")
 * x86-64 + gcc only: the double->integer casts below rely on the same cvttsd2si behaviour as the
 * reference build.
 */
#include "sz_oracle.h"
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define x86_d2u64(d) ((uint64_t)(d))
#define x86_d2i64(d) ((int64_t)(d))
#define x86_f2u64(f) ((uint64_t)(f))

/* ---- byte helpers: ByteToolkit.c:198 (int BE), :278 (long BE), :442 (float BE), :461 (double BE) ---- */
static void put_be32(unsigned char *b, uint32_t v) { b[0] = v >> 24; b[1] = v >> 16; b[2] = v >> 8; b[3] = v; }
static void put_be16(unsigned char *b, uint16_t v) { b[0] = v >> 8; b[1] = v; }
static void put_be64(unsigned char *b, uint64_t v) { for (int i = 0; i < 8; i++) b[i] = (unsigned char)(v >> (56 - 8 * i)); }
static uint32_t get_be32(const unsigned char *b) { return ((uint32_t)b[0] << 24) | ((uint32_t)b[1] << 16) | ((uint32_t)b[2] << 8) | b[3]; }
static uint64_t get_be64(const unsigned char *b) { uint64_t v = 0; for (int i = 0; i < 8; i++) v = (v << 8) | b[i]; return v; }
static void put_be_f(unsigned char *b, float f) { uint32_t u; memcpy(&u, &f, 4); put_be32(b, u); }
static void put_be_d(unsigned char *b, double f) { uint64_t u; memcpy(&u, &f, 8); put_be64(b, u); }
static float get_be_f(const unsigned char *b) { uint32_t u = get_be32(b); float f; memcpy(&f, &u, 4); return f; }
static double get_be_d(const unsigned char *b) { uint64_t u = get_be64(b); double f; memcpy(&f, &u, 8); return f; }

/* conf.c:35-46 */
static unsigned int round_up_pow2(unsigned int base)
{
    base -= 1;
    base |= base >> 1; base |= base >> 2; base |= base >> 4; base |= base >> 8; base |= base >> 16;
    return base + 1;
}

/* sz.c:117-160 */
static size_t data_length(size_t r5, size_t r4, size_t r3, size_t r2, size_t r1)
{
    if (r1 == 0) return 0;
    if (r2 == 0) return r1;
    if (r3 == 0) return r1 * r2;
    if (r4 == 0) return r1 * r2 * r3;
    if (r5 == 0) return r1 * r2 * r3 * r4;
    return r1 * r2 * r3 * r4 * r5;
}

/* filterDimension, sz.c:162-282: every dimension of extent 1 is removed, order kept */
int szo_filter_dims(size_t r5, size_t r4, size_t r3, size_t r2, size_t r1, size_t *r)
{
    size_t in[5] = {r1, r2, r3, r4, r5};
    int dim = 0, changed = 0, n = 0;
    while (dim < 5 && in[dim] != 0) dim++;
    for (int i = 0; i < 5; i++) r[i] = 0;
    if (dim == 1) { r[0] = r1; return (r1 < 1) ? 2 : 0; }
    for (int i = 0; i < dim; i++) { if (in[i] == 1) changed = 1; else r[n++] = in[i]; }
    return changed;
}

/* block geometry per axis: sz.h:93-123 (SZ_COMPUTE_*_NUMBER_OF_BLOCKS, SZ_COMPUTE_BLOCKCOUNT) */
typedef struct szo_axis { size_t n, nb, early, late, split; } szo_axis;
static void axis_init(szo_axis *a, size_t n, size_t B)
{
    a->n = n;
    a->nb = (n <= B) ? 1 : n / B;
    a->early = a->late = n / a->nb;
    a->split = n % a->nb;
    if (a->split) a->early++;
}
static size_t axis_edge(const szo_axis *a, size_t b) { return b < a->split ? a->early : a->late; }
static size_t axis_off(const szo_axis *a, size_t b) { return b < a->split ? b * a->early : b * a->late + a->split; }

/* ---- stream header: dataCompression.c:686-709 + ByteToolkit.c:874-972 ------------------------------ */
typedef struct szo_hdr { const szo_config *cfg; int mode, dtype; double abs, rel, vmin, vmax; } szo_hdr;

static size_t write_header(unsigned char *p, const szo_hdr *h, int flagbits, int mdlen)
{
    const szo_config *c = h->cfg;
    p[0] = 2; p[1] = 1; p[2] = 12;                       /* versionNumber, defines.h:14-16 */
    int same = flagbits | 0x40;                           /* SZ_SIZE_TYPE == 8 */
    if (c->protectValueRange) same |= 0x04;
    p[3] = (unsigned char)same;
    unsigned char *m = p + 4;
    memset(m, 0, (size_t)mdlen);
    int optQuant = (c->quantization_intervals == 0);
    int gz = 0;
    if (c->gzipMode == 1) gz = 0; else if (c->gzipMode == 0) gz = 1; else if (c->gzipMode == 9) gz = 2;
    unsigned char buf = (unsigned char)optQuant;
    buf = (unsigned char)((buf << 1) | c->dataEndianType);
    buf = (unsigned char)((buf << 1) | 0 /* sysEndianType little */);
    buf = (unsigned char)((buf << 2) | c->szMode);
    buf = (unsigned char)((buf << 2) | gz);
    m[0] = buf;
    put_be16(m + 1, (uint16_t)(int16_t)c->sampleDistance);
    short t2 = (short)(c->predThreshold * 10000);
    put_be16(m + 3, (uint16_t)t2);
    m[5] = (unsigned char)((h->mode << 4) | (h->dtype & 0x17));
    if (h->mode == SZO_ABS) put_be_f(m + 6, (float)h->abs);
    else if (h->mode == SZO_REL) put_be_f(m + 10, (float)h->rel);
    else { put_be_f(m + 6, (float)h->abs); put_be_f(m + 10, (float)h->rel); }
    m[14] = (unsigned char)c->sol_ID;
    put_be32(m + 16, optQuant ? c->max_quant_intervals : c->quantization_intervals);
    if (h->dtype == SZO_FLOAT) { put_be_f(m + 20, (float)h->vmin); put_be_f(m + 24, (float)h->vmax); }
    else { put_be_d(m + 20, h->vmin); put_be_d(m + 28, h->vmax); }
    return (size_t)(4 + mdlen);
}

/* ---- Huffman: Huffman.c:17-185 (tree), :122-157 (codes), :205-308 (encode), :310-343 (decode),
 *               :443-585 (serialise), :656-788 (reconstruct) ---------------------------------------- */
typedef struct szo_hnode { size_t freq; int left, right; unsigned int c; unsigned char t; } szo_hnode;
typedef struct szo_huff {
    int stateNum; szo_hnode *pool; int n_nodes; int *qq; int qend;
    uint64_t *code; unsigned char *clen; int root; unsigned int nodeCount; int overflow;
} szo_huff;

static void huff_init(szo_huff *h, int stateNum)
{
    memset(h, 0, sizeof(*h));
    h->stateNum = stateNum;
    h->pool = (szo_hnode *)calloc((size_t)stateNum * 2 + 2, sizeof(szo_hnode));
    h->qq = (int *)calloc((size_t)stateNum * 2 + 2, sizeof(int));
    h->code = (uint64_t *)calloc((size_t)stateNum, sizeof(uint64_t));
    h->clen = (unsigned char *)calloc((size_t)stateNum, 1);
    h->qend = 1; h->root = -1;
}
static void huff_free(szo_huff *h) { free(h->pool); free(h->qq); free(h->code); free(h->clen); memset(h, 0, sizeof(*h)); }

static void h_qinsert(szo_huff *h, int n)               /* Huffman.c:76-85 */
{
    int j, i = h->qend++;
    while ((j = i >> 1)) {
        if (h->pool[h->qq[j]].freq <= h->pool[n].freq) break;
        h->qq[i] = h->qq[j]; i = j;
    }
    h->qq[i] = n;
}
static int h_qremove(szo_huff *h)                        /* Huffman.c:87-114 */
{
    int i = 1, l, n = h->qq[1];
    if (h->qend < 2) return -1;
    h->qend--;
    h->qq[1] = h->qq[h->qend];
    while ((l = i << 1) < h->qend) {
        if (l + 1 < h->qend && h->pool[h->qq[l + 1]].freq < h->pool[h->qq[l]].freq) l++;
        if (h->pool[h->qq[i]].freq > h->pool[h->qq[l]].freq) { int p = h->qq[i]; h->qq[i] = h->qq[l]; h->qq[l] = p; i = l; }
        else break;
    }
    return n;
}
static void h_build_code(szo_huff *h, int n, int len, uint64_t bits)   /* Huffman.c:122-157, len<=64 only */
{
    szo_hnode *nd = &h->pool[n];
    if (nd->t) {
        if (len > 64) { h->overflow = 1; return; }
        h->code[nd->c] = len ? (bits << (64 - len)) : 0;
        h->clen[nd->c] = (unsigned char)len;
        return;
    }
    if (len >= 64) { h->overflow = 1; return; }
    h_build_code(h, nd->left, len + 1, bits << 1);
    h_build_code(h, nd->right, len + 1, (bits << 1) | 1);
}
static int huff_build(szo_huff *h, const int *s, size_t n)            /* init(), Huffman.c:165-185 */
{
    size_t *freq = (size_t *)calloc((size_t)h->stateNum * 2, sizeof(size_t));
    for (size_t i = 0; i < n; i++) freq[s[i]]++;
    unsigned int leaves = 0;
    for (int i = 0; i < h->stateNum * 2; i++) if (freq[i]) {
        int id = h->n_nodes++;
        h->pool[id].c = (unsigned int)i; h->pool[id].freq = freq[i]; h->pool[id].t = 1; h->pool[id].left = h->pool[id].right = -1;
        h_qinsert(h, id); leaves++;
    }
    free(freq);
    while (h->qend > 2) {
        /* gcc evaluates the two qremove() arguments right-to-left (Huffman.c:181): first removed -> right */
        int b = h_qremove(h);
        int a = h_qremove(h);
        int id = h->n_nodes++;
        h->pool[id].left = a; h->pool[id].right = b; h->pool[id].freq = h->pool[a].freq + h->pool[b].freq; h->pool[id].t = 0; h->pool[id].c = 0;
        h_qinsert(h, id);
    }
    h->root = h->qq[1];
    h_build_code(h, h->root, 0, 0);
    h->nodeCount = leaves * 2 - 1;
    return h->overflow ? -1 : 0;
}

typedef struct pad_ctx { uint32_t *L, *R, *C; unsigned char *t; unsigned int n_inode; } pad_ctx;
static void h_pad(szo_huff *h, pad_ctx *c, unsigned int i, int root)   /* pad_tree_*, Huffman.c:443-501 */
{
    c->C[i] = h->pool[root].c; c->t[i] = h->pool[root].t;
    if (!h->pool[root].t) {
        int l = h->pool[root].left, r = h->pool[root].right;
        if (l >= 0) { c->n_inode++; c->L[i] = c->n_inode; h_pad(h, c, c->n_inode, l); }
        if (r >= 0) { c->n_inode++; c->R[i] = c->n_inode; h_pad(h, c, c->n_inode, r); }
    }
}
static unsigned int huff_serialize(szo_huff *h, unsigned char **out)   /* convert_HuffTree_to_bytes_anyStates, Huffman.c:503-585 */
{
    unsigned int n = h->nodeCount;
    pad_ctx c; c.n_inode = 0;
    c.L = (uint32_t *)calloc(n, 4); c.R = (uint32_t *)calloc(n, 4); c.C = (uint32_t *)calloc(n, 4); c.t = (unsigned char *)calloc(n, 1);
    h_pad(h, &c, 0, h->root);
    int w = (n <= 256) ? 1 : (n <= 65536) ? 2 : 4;
    unsigned int total = 1 + 2 * w * n + 4 * n + n;
    unsigned char *o = (unsigned char *)malloc(total);
    o[0] = 0;                                            /* sysEndianType little */
    unsigned char *pl = o + 1, *pr = o + 1 + (size_t)w * n, *pc = o + 1 + 2 * (size_t)w * n, *pt = pc + 4 * (size_t)n;
    for (unsigned int i = 0; i < n; i++) {
        if (w == 1) { pl[i] = (unsigned char)c.L[i]; pr[i] = (unsigned char)c.R[i]; }
        else if (w == 2) { uint16_t a = (uint16_t)c.L[i], b = (uint16_t)c.R[i]; memcpy(pl + 2 * i, &a, 2); memcpy(pr + 2 * i, &b, 2); }
        else { memcpy(pl + 4 * i, &c.L[i], 4); memcpy(pr + 4 * i, &c.R[i], 4); }
        memcpy(pc + 4 * i, &c.C[i], 4);
        pt[i] = c.t[i];
    }
    free(c.L); free(c.R); free(c.C); free(c.t);
    *out = o;
    return total;
}
/* encode(), Huffman.c:205-308, restated as an MSB-first bit writer (identical bytes for code lengths <= 64) */
static size_t huff_encode_tbl(const uint64_t *code, const unsigned char *clen, const int *s, size_t n, unsigned char *out)
{
    uint64_t acc = 0; int nacc = 0; size_t o = 0;
    for (size_t i = 0; i < n; i++) {
        int len = clen[s[i]]; uint64_t c = code[s[i]];
        int pos = 0;
        while (pos < len) {
            int take = 64 - nacc; if (take > len - pos) take = len - pos;
            uint64_t chunk = (take == 64) ? c : ((c << pos) >> (64 - take));
            acc = (take == 64) ? chunk : ((acc << take) | chunk);
            nacc += take; pos += take;
            if (nacc == 64) { put_be64(out + o, acc); o += 8; acc = 0; nacc = 0; }
        }
    }
    if (nacc) { acc <<= (64 - nacc); for (int b = 0; b < (nacc + 7) / 8; b++) out[o++] = (unsigned char)(acc >> (56 - 8 * b)); }
    return o;
}
static size_t huff_encode(szo_huff *h, const int *s, size_t n, unsigned char *out) { return huff_encode_tbl(h->code, h->clen, s, n, out); }

/* reconstruct_HuffTree_from_bytes_anyStates (Huffman.c:656-788) + decode (:310-343) */
static int huff_decode_stream(const unsigned char *tb, unsigned int n, const unsigned char *bits, size_t count, int *out)
{
    int w = (n <= 256) ? 1 : (n <= 65536) ? 2 : 4;
    const unsigned char *pl = tb + 1, *pr = tb + 1 + (size_t)w * n, *pc = tb + 1 + 2 * (size_t)w * n, *pt = pc + 4 * (size_t)n;
#define LV(base, i) (w == 1 ? (uint32_t)(base)[i] : w == 2 ? (uint32_t)((base)[2 * (i)] | ((base)[2 * (i) + 1] << 8)) : \
                     (uint32_t)((base)[4 * (i)] | ((base)[4 * (i) + 1] << 8) | ((base)[4 * (i) + 2] << 16) | ((uint32_t)(base)[4 * (i) + 3] << 24)))
    if (pt[0]) { uint32_t c; memcpy(&c, pc, 4); for (size_t i = 0; i < count; i++) out[i] = (int)c; return 0; }
    uint32_t node = 0; size_t cnt = 0;
    for (size_t i = 0; cnt < count; i++) {
        int bit = (bits[i >> 3] >> (7 - (i & 7))) & 1;
        node = bit ? LV(pr, node) : LV(pl, node);
        if (node == 0 || node >= n) return -1;
        if (pt[node]) { uint32_t c; memcpy(&c, pc + 4 * (size_t)node, 4); out[cnt++] = (int)c; node = 0; }
    }
#undef LV
    return 0;
}

int szo_huffman_build(const int *s, size_t n, int stateNum, unsigned char **treeBytes, unsigned int *treeByteSize,
                      unsigned int *nodeCount, uint64_t *code, unsigned char *clen)
{
    szo_huff h; huff_init(&h, stateNum);
    int rc = huff_build(&h, s, n);
    if (rc == 0) {
        *treeByteSize = huff_serialize(&h, treeBytes); *nodeCount = h.nodeCount;
        memcpy(code, h.code, sizeof(uint64_t) * (size_t)stateNum); memcpy(clen, h.clen, (size_t)stateNum);
    }
    huff_free(&h);
    return rc;
}
size_t szo_huffman_encode(const int *s, size_t n, const uint64_t *code, const unsigned char *clen, unsigned char *out)
{ return huff_encode_tbl(code, clen, s, n, out); }
int szo_huffman_decode(const unsigned char *treeBytes, unsigned int nodeCount, const unsigned char *bits, size_t n, int *out)
{ return huff_decode_stream(treeBytes, nodeCount, bits, n, out); }

/* ---- type-generic part ---------------------------------------------------------------------------- */
#define CAT_(a, b) a##b
#define CAT(a, b) CAT_(a, b)
#define FN(name) CAT(name, SUF)

#define T float
#define SUF _f
#define IS_FLOAT 1
#define MDLEN 28
#define FABS_T(v) fabsf(v)
#include "sz_oracle_tmpl.h"
#undef T
#undef SUF
#undef IS_FLOAT
#undef MDLEN
#undef FABS_T

#define T double
#define SUF _d
#define IS_FLOAT 0
#define MDLEN 36
#define FABS_T(v) fabs(v)
#include "sz_oracle_tmpl.h"
#undef T
#undef SUF
#undef IS_FLOAT
#undef MDLEN
#undef FABS_T

void szo_default_config(szo_config *c)
{
    memset(c, 0, sizeof(*c));
    c->max_quant_intervals = 65536; c->quantization_intervals = 0; c->sampleDistance = 100; c->predThreshold = 0.99f;
    c->szMode = 0; c->gzipMode = 3; c->protectValueRange = 0; c->sol_ID = 101; c->dataEndianType = 0;
    c->conf_absErrBound = 1e-4; c->conf_relBoundRatio = 1e-4; c->psnr = 80; c->normErr = 0.05;
}

int szo_compress(int dtype, const void *data, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1,
                 int mode, double absb, double relb, const szo_config *cfg, unsigned char **out, size_t *outSize, szo_trace *tr)
{
    size_t r[5];
    szo_filter_dims(r5, r4, r3, r2, r1, r);
    if (dtype == SZO_FLOAT) return compress_args_f((const float *)data, r, mode, absb, relb, cfg, out, outSize, tr);
    if (dtype == SZO_DOUBLE) return compress_args_d((const double *)data, r, mode, absb, relb, cfg, out, outSize, tr);
    return -3;
}

int szo_decompress(int dtype, const unsigned char *bytes, size_t len, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1, void **out)
{
    size_t r[5];
    szo_filter_dims(r5, r4, r3, r2, r1, r);
    if (dtype == SZO_FLOAT) return decompress_args_f(bytes, len, r, out);
    if (dtype == SZO_DOUBLE) return decompress_args_d(bytes, len, r, out);
    return -3;
}

void szo_trace_free(szo_trace *t)
{
    if (!t) return;
    free(t->reg_params); free(t->indicator); free(t->coeff_type); free(t->coeff_last); free(t->type); free(t->unpred); free(t->recon);
    memset(t, 0, sizeof(*t));
}
void szo_free(void *p) { free(p); }
