/*
 * sz_host.c -- the C host layer of libSZ_b200.so: SZ's public API (include/sz.h) for the float/double
 * blocked-regression path, on top of the CUDA shim (include/szb200.h).
 *
 * Mirrors the reference's L3 (sz/src/sz.c, sz/src/conf.c) and the host-side orchestration of its L2
 * (SZ_compress_args_float, sz/src/sz_float.c:2811-3043; SZ_decompress_args_float, sz/src/szd_float.c:50-183):
 * dimension clean-up, range -> error bound, the constant / lossless / tiny-input containers, header bytes.
 * Everything data-parallel is delegated to kernels; nothing here touches the bulk data on the CPU and there is
 * no CPU fallback: if the shim reports an error the call fails (NULL / SZ_NSCS).
 */
#define _GNU_SOURCE
#include <ctype.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#include <math.h>
#include <pthread.h>
#include <time.h>
#include <sys/mman.h>
/* Buffers handed back to the caller (free()-able).  Large ones are 2 MiB-aligned and marked for transparent huge pages: a
 * device-to-host copy into FRESH 4 KiB pages is bound by page faults (measured 2 GB/s); with huge pages it is not. */
static void *host_alloc_result(size_t n) {
    const size_t HP = (size_t)2 << 20;
    if (n >= 2 * HP) {
        void *p = NULL;
        const size_t len = (n + HP - 1) & ~(HP - 1);
        if (posix_memalign(&p, HP, len) == 0 && p) {
#ifdef MADV_HUGEPAGE
            (void)madvise(p, len, MADV_HUGEPAGE);
#endif
            return p;
        }
    }
    return malloc(n ? n : 1);
}
/* The result buffer of the host-pointer compress call is allocated BEFORE the GPU work, from the size of the previous stream,
 * and a helper thread touches its pages while the GPU is busy; the final copy then lands on resident pages.  A wrong guess
 * only costs a second allocation. */
typedef struct prefault_job { unsigned char *p; size_t n; } prefault_job;
static void *prefault_thread(void *arg) {
    prefault_job *j = (prefault_job *)arg;
#ifdef MADV_POPULATE_WRITE
    if (madvise(j->p, j->n, MADV_POPULATE_WRITE) == 0) return NULL;
#endif
    for (size_t o = 0; o < j->n; o += 4096) ((volatile unsigned char *)j->p)[o] = 0;
    return NULL;
}
static size_t g_last_stream_bytes = 0;                /* guarded by g_api_mu like the rest of the global state */
static double now_ms(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec * 1e3 + t.tv_nsec * 1e-6; }
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include "sz.h"
#include "szb200.h"

/* ---- exported data globals (ref: sz/src/sz.c:33-53) ---------------------------------------------------- */
int versionNumber[4] = {SZ_VER_MAJOR, SZ_VER_MINOR, SZ_VER_BUILD, SZ_VER_REVISION};
int dataEndianType = LITTLE_ENDIAN_DATA;
int sysEndianType = LITTLE_ENDIAN_SYSTEM;
sz_params *confparams_cpr = NULL;
sz_params *confparams_dec = NULL;
sz_exedata *exe_params = NULL;
/* The batch entry points run several jobs at once on worker threads.  Every worker owns private copies of the three parameter
 * blocks and publishes them through these thread-local overrides; the single-call API leaves them NULL and works on the
 * exported globals exactly like the reference. */
static __thread sz_params *tl_cpr = NULL, *tl_dec = NULL;
static __thread sz_exedata *tl_exe = NULL;
static inline sz_params *CPR(void) { return tl_cpr ? tl_cpr : confparams_cpr; }
static inline sz_params *DEC(void) { return tl_dec ? tl_dec : confparams_dec; }
static inline sz_exedata *EXE(void) { return tl_exe ? tl_exe : exe_params; }


static pthread_mutex_t g_api_mu = PTHREAD_MUTEX_INITIALIZER;   /* the globals above are shared state */

/* ---- byte helpers (ref: sz/src/ByteToolkit.c:198, :278, :442, :461, :817, :827) ------------------------------ */
static void put_be16(unsigned char *b, uint16_t v) { b[0] = (unsigned char)(v >> 8); b[1] = (unsigned char)v; }
static void put_be32(unsigned char *b, uint32_t v) { b[0] = (unsigned char)(v >> 24); b[1] = (unsigned char)(v >> 16); b[2] = (unsigned char)(v >> 8); b[3] = (unsigned char)v; }
static void put_be64(unsigned char *b, uint64_t v) { for (int i = 0; i < 8; i++) b[i] = (unsigned char)(v >> (56 - 8 * i)); }
static uint16_t get_be16(const unsigned char *b) { return (uint16_t)((b[0] << 8) | b[1]); }
static uint32_t get_be32(const unsigned char *b) { return ((uint32_t)b[0] << 24) | ((uint32_t)b[1] << 16) | ((uint32_t)b[2] << 8) | b[3]; }
static uint64_t get_be64(const unsigned char *b) { uint64_t v = 0; for (int i = 0; i < 8; i++) v = (v << 8) | b[i]; return v; }
static void put_be_float(unsigned char *b, float f) { uint32_t u; memcpy(&u, &f, 4); put_be32(b, u); }
static void put_be_double(unsigned char *b, double f) { uint64_t u; memcpy(&u, &f, 8); put_be64(b, u); }
static float get_be_float(const unsigned char *b) { uint32_t u = get_be32(b); float f; memcpy(&f, &u, 4); return f; }
static double get_be_double(const unsigned char *b) { uint64_t u = get_be64(b); double f; memcpy(&f, &u, 8); return f; }

static void detect_endianness(void) { int x = 1; sysEndianType = (*(char *)&x == 1) ? LITTLE_ENDIAN_SYSTEM : BIG_ENDIAN_SYSTEM; }

/* ---- a minimal sz.config reader: "[SECTION]" / "key = value" lines, '#' and ';' comments.
 * Key names, defaults and error behaviour follow SZ_ReadConf (ref: sz/src/conf.c:74-400). -------------------- */
#define CFG_MAX 64
typedef struct { char key[96]; char val[160]; } cfg_kv;
typedef struct { cfg_kv kv[CFG_MAX]; int n; } cfg_t;

static void str_trim(char *s) {
    size_t n = strlen(s);
    while (n && isspace((unsigned char)s[n - 1])) s[--n] = 0;
    size_t i = 0; while (s[i] && isspace((unsigned char)s[i])) i++;
    if (i) memmove(s, s + i, n - i + 1);
}
static void str_lower(char *s) { for (; *s; s++) *s = (char)tolower((unsigned char)*s); }

static int cfg_load(const char *path, cfg_t *c) {
    FILE *f = fopen(path, "r");
    if (!f) return -1;
    char line[512], section[24] = "";
    c->n = 0;
    while (fgets(line, sizeof(line), f)) {
        str_trim(line);
        if (!line[0] || line[0] == '#' || line[0] == ';') continue;
        if (line[0] == '[') { char *e = strchr(line, ']'); if (e) { *e = 0; snprintf(section, sizeof(section), "%.23s", line + 1); str_trim(section); str_lower(section); } continue; }
        char *eq = strchr(line, '=');
        if (!eq || c->n >= CFG_MAX) continue;
        *eq = 0;
        char key[30], val[160];
        snprintf(key, sizeof(key), "%.29s", line); str_trim(key); str_lower(key);
        snprintf(val, sizeof(val), "%s", eq + 1);
        char *cm = strpbrk(val, ";#"); if (cm) *cm = 0;
        str_trim(val);
        size_t vl = strlen(val);
        if (vl >= 2 && ((val[0] == '"' && val[vl - 1] == '"') || (val[0] == '\'' && val[vl - 1] == '\''))) { val[vl - 1] = 0; memmove(val, val + 1, vl - 1); }
        snprintf(c->kv[c->n].key, sizeof(c->kv[c->n].key), "%s:%s", section, key);
        snprintf(c->kv[c->n].val, sizeof(c->kv[c->n].val), "%s", val);
        c->n++;
    }
    fclose(f);
    return 0;
}
static const char *cfg_str(const cfg_t *c, const char *key, const char *def) {
    char k[96]; snprintf(k, sizeof(k), "%s", key); str_lower(k);
    for (int i = 0; i < c->n; i++) if (strcmp(c->kv[i].key, k) == 0) return c->kv[i].val;
    return def;
}
static int cfg_int(const cfg_t *c, const char *key, int def) { const char *s = cfg_str(c, key, NULL); return s ? (int)strtol(s, NULL, 0) : def; }
static double cfg_dbl(const cfg_t *c, const char *key, double def) { const char *s = cfg_str(c, key, NULL); return s ? atof(s) : def; }

/* ref: sz/src/conf.c:48-52 */
static void updateQuantizationInfo(int quant_intervals) { exe_params->intvCapacity = quant_intervals; exe_params->intvRadius = quant_intervals / 2; }

/* ref: sz/src/conf.c:74-400 (SZ_ReadConf) */
static int sz_read_conf(const char *path) {
    confparams_cpr = (sz_params *)calloc(1, sizeof(sz_params));
    exe_params = (sz_exedata *)calloc(1, sizeof(sz_exedata));
    detect_endianness();
    confparams_cpr->plus_bits = 3;
    if (path == NULL) {                              /* built-in defaults, ref: conf.c:99-141 */
        dataEndianType = LITTLE_ENDIAN_DATA;
        confparams_cpr->sol_ID = SZ;
        confparams_cpr->max_quant_intervals = 65536;
        confparams_cpr->maxRangeRadius = confparams_cpr->max_quant_intervals / 2;
        exe_params->intvCapacity = (int)confparams_cpr->maxRangeRadius * 2;
        exe_params->intvRadius = (int)confparams_cpr->maxRangeRadius;
        confparams_cpr->quantization_intervals = 0;
        exe_params->optQuantMode = 1;
        confparams_cpr->predThreshold = 0.99f;
        confparams_cpr->sampleDistance = 100;
        confparams_cpr->szMode = SZ_BEST_COMPRESSION;
        confparams_cpr->losslessCompressor = ZSTD_COMPRESSOR;
        confparams_cpr->gzipMode = 3;
        confparams_cpr->errorBoundMode = PSNR;
        confparams_cpr->psnr = 90;
        confparams_cpr->absErrBound = 1E-4;
        confparams_cpr->relBoundRatio = 1E-4;
        confparams_cpr->accelerate_pw_rel_compression = 1;
        confparams_cpr->pw_relBoundRatio = 1E-3;
        confparams_cpr->segment_size = 36;
        confparams_cpr->pwr_type = 0;
        confparams_cpr->snapshotCmprStep = 5;
        confparams_cpr->withRegression = SZ_WITH_LINEAR_REGRESSION;
        confparams_cpr->randomAccess = 0;
        confparams_cpr->protectValueRange = 0;
        return SZ_SCES;
    }
    if (access(path, F_OK) != 0) { printf("[SZ] Configuration file NOT accessible.\n"); return SZ_NSCS; }
    cfg_t c;
    if (cfg_load(path, &c) != 0) { printf("[SZ] Iniparser failed to parse the conf. file.\n"); return SZ_NSCS; }
    const char *s = cfg_str(&c, "ENV:dataEndianType", "LITTLE_ENDIAN_DATA");
    if (strcmp(s, "LITTLE_ENDIAN_DATA") == 0) dataEndianType = LITTLE_ENDIAN_DATA;
    else if (strcmp(s, "BIG_ENDIAN_DATA") == 0) dataEndianType = BIG_ENDIAN_DATA;
    else { printf("Error: Wrong dataEndianType: please set it correctly in sz.config.\n"); return SZ_NSCS; }
    s = cfg_str(&c, "ENV:sol_name", "(null)");
    if (strcmp(s, "SZ") == 0) confparams_cpr->sol_ID = SZ;
    else if (strcmp(s, "SZ_Transpose") == 0) confparams_cpr->sol_ID = SZ_Transpose;
    else { printf("[SZ] Error: wrong solution name (please check sz.config file), sol=%s\n", s); return SZ_NSCS; }

    int max_quant_intervals = cfg_int(&c, "PARAMETER:max_quant_intervals", 65536);
    confparams_cpr->max_quant_intervals = (unsigned int)max_quant_intervals;
    int quantization_intervals = cfg_int(&c, "PARAMETER:quantization_intervals", 0);
    confparams_cpr->quantization_intervals = (unsigned int)quantization_intervals;
    if (quantization_intervals > 0) {
        updateQuantizationInfo(quantization_intervals);
        confparams_cpr->max_quant_intervals = (unsigned int)quantization_intervals;
        exe_params->optQuantMode = 0;
    } else {
        confparams_cpr->maxRangeRadius = (unsigned int)max_quant_intervals / 2;
        exe_params->intvCapacity = (int)confparams_cpr->maxRangeRadius * 2;
        exe_params->intvRadius = (int)confparams_cpr->maxRangeRadius;
        exe_params->optQuantMode = 1;
    }
    if (quantization_intervals % 2 != 0) { printf("Error: quantization_intervals must be an even number!\n"); return SZ_NSCS; }
    confparams_cpr->predThreshold = (float)cfg_dbl(&c, "PARAMETER:predThreshold", 0);
    confparams_cpr->sampleDistance = cfg_int(&c, "PARAMETER:sampleDistance", 0);
    s = cfg_str(&c, "PARAMETER:szMode", NULL);
    if (s == NULL) { printf("[SZ] Error: Null szMode setting (please check sz.config file)\n"); return SZ_NSCS; }
    else if (strcmp(s, "SZ_BEST_SPEED") == 0) confparams_cpr->szMode = SZ_BEST_SPEED;
    else if (strcmp(s, "SZ_DEFAULT_COMPRESSION") == 0) confparams_cpr->szMode = SZ_DEFAULT_COMPRESSION;
    else if (strcmp(s, "SZ_BEST_COMPRESSION") == 0) confparams_cpr->szMode = SZ_BEST_COMPRESSION;
    else { printf("[SZ] Error: Wrong szMode setting (please check sz.config file)\n"); return SZ_NSCS; }
    s = cfg_str(&c, "PARAMETER:losslessCompressor", "ZSTD_COMPRESSOR");
    if (strcmp(s, "GZIP_COMPRESSOR") == 0) confparams_cpr->losslessCompressor = GZIP_COMPRESSOR;
    else if (strcmp(s, "ZSTD_COMPRESSOR") == 0) confparams_cpr->losslessCompressor = ZSTD_COMPRESSOR;
    else { printf("[SZ] Error: Wrong losslessCompressor setting (please check sz.config file)\nNo Such a lossless compressor: %s\n", s); return SZ_NSCS; }
    s = cfg_str(&c, "PARAMETER:withLinearRegression", "YES");
    confparams_cpr->withRegression = (strcmp(s, "YES") == 0 || strcmp(s, "yes") == 0) ? SZ_WITH_LINEAR_REGRESSION : SZ_NO_REGRESSION;
    s = cfg_str(&c, "PARAMETER:gzipMode", "Gzip_BEST_SPEED");
    if (strcmp(s, "Gzip_NO_COMPRESSION") == 0) confparams_cpr->gzipMode = 0;
    else if (strcmp(s, "Gzip_BEST_SPEED") == 0) confparams_cpr->gzipMode = 1;
    else if (strcmp(s, "Gzip_BEST_COMPRESSION") == 0) confparams_cpr->gzipMode = 9;
    else if (strcmp(s, "Gzip_DEFAULT_COMPRESSION") == 0) confparams_cpr->gzipMode = -1;
    else { printf("[SZ] Error: Wrong gzip Mode (please check sz.config file)\n"); return SZ_NSCS; }
    s = cfg_str(&c, "PARAMETER:zstdMode", "Zstd_HIGH_SPEED");   /* overrides gzipMode, as in the reference (conf.c:275-296) */
    if (strcmp(s, "Zstd_BEST_SPEED") == 0) confparams_cpr->gzipMode = 1;
    else if (strcmp(s, "Zstd_HIGH_SPEED") == 0) confparams_cpr->gzipMode = 3;
    else if (strcmp(s, "Zstd_HIGH_COMPRESSION") == 0) confparams_cpr->gzipMode = 19;
    else if (strcmp(s, "Zstd_BEST_COMPRESSION") == 0) confparams_cpr->gzipMode = 22;
    else if (strcmp(s, "Zstd_DEFAULT_COMPRESSION") == 0) confparams_cpr->gzipMode = 3;
    else { printf("[SZ] Error: Wrong zstd Mode (please check sz.config file)\n"); return SZ_NSCS; }
    s = cfg_str(&c, "PARAMETER:protectValueRange", "YES");
    confparams_cpr->protectValueRange = strcmp(s, "YES") == 0 ? 1 : 0;
    confparams_cpr->randomAccess = cfg_int(&c, "PARAMETER:randomAccess", 0);
    confparams_cpr->snapshotCmprStep = cfg_int(&c, "PARAMETER:snapshotCmprStep", 5);
    s = cfg_str(&c, "PARAMETER:errorBoundMode", NULL);
    if (s == NULL) { printf("[SZ] Error: Null error bound setting (please check sz.config file)\n"); return SZ_NSCS; }
    else if (!strcmp(s, "ABS") || !strcmp(s, "abs")) confparams_cpr->errorBoundMode = ABS;
    else if (!strcmp(s, "REL") || !strcmp(s, "rel") || !strcmp(s, "VR_REL") || !strcmp(s, "vr_rel")) confparams_cpr->errorBoundMode = REL;
    else if (!strcmp(s, "ABS_AND_REL") || !strcmp(s, "abs_and_rel")) confparams_cpr->errorBoundMode = ABS_AND_REL;
    else if (!strcmp(s, "ABS_OR_REL") || !strcmp(s, "abs_or_rel")) confparams_cpr->errorBoundMode = ABS_OR_REL;
    else if (!strcmp(s, "PW_REL") || !strcmp(s, "pw_rel")) confparams_cpr->errorBoundMode = PW_REL;
    else if (!strcmp(s, "PSNR") || !strcmp(s, "psnr")) confparams_cpr->errorBoundMode = PSNR;
    else if (!strcmp(s, "ABS_AND_PW_REL") || !strcmp(s, "abs_and_pw_rel")) confparams_cpr->errorBoundMode = ABS_AND_PW_REL;
    else if (!strcmp(s, "ABS_OR_PW_REL") || !strcmp(s, "abs_or_pw_rel")) confparams_cpr->errorBoundMode = ABS_OR_PW_REL;
    else if (!strcmp(s, "REL_AND_PW_REL") || !strcmp(s, "rel_and_pw_rel")) confparams_cpr->errorBoundMode = REL_AND_PW_REL;
    else if (!strcmp(s, "REL_OR_PW_REL") || !strcmp(s, "rel_or_pw_rel")) confparams_cpr->errorBoundMode = REL_OR_PW_REL;
    else if (!strcmp(s, "NORM") || !strcmp(s, "norm")) confparams_cpr->errorBoundMode = NORM;
    else { printf("[SZ] Error: Wrong error bound mode (please check sz.config file)\n"); return SZ_NSCS; }
    confparams_cpr->absErrBound = cfg_dbl(&c, "PARAMETER:absErrBound", 0);
    confparams_cpr->relBoundRatio = cfg_dbl(&c, "PARAMETER:relBoundRatio", 0);
    confparams_cpr->psnr = cfg_dbl(&c, "PARAMETER:psnr", 0);
    confparams_cpr->normErr = cfg_dbl(&c, "PARAMETER:normErr", 0);
    confparams_cpr->pw_relBoundRatio = cfg_dbl(&c, "PARAMETER:pw_relBoundRatio", 0);
    confparams_cpr->segment_size = cfg_int(&c, "PARAMETER:segment_size", 0);
    confparams_cpr->accelerate_pw_rel_compression = cfg_int(&c, "PARAMETER:accelerate_pw_rel_compression", 1);
    s = cfg_str(&c, "PARAMETER:pwr_type", "MIN");
    confparams_cpr->pwr_type = strcmp(s, "AVG") == 0 ? 1 : (strcmp(s, "MAX") == 0 ? 2 : 0);
    return SZ_SCES;
}

/* ref: sz/src/sz.c:60-73 */
int SZ_Init(const char *configFilePath) {
    if (sz_read_conf(configFilePath) == SZ_NSCS) return SZ_NSCS;
    exe_params->SZ_SIZE_TYPE = sizeof(size_t);
    /* The reference's built-in configuration (SZ_Init(NULL), conf.c:114) asks for SZ_BEST_COMPRESSION, i.e. the zstd pass this
     * library does not have, so a caller that never sets szMode (example/test_threadsafe.c) would only get refusals.
     * SZB200_SZMODE=SZ_BEST_SPEED in the environment overrides the configured mode -- an explicit, outside decision. */
    const char *m = getenv("SZB200_SZMODE");
    if (m && strcmp(m, "SZ_BEST_SPEED") == 0) confparams_cpr->szMode = SZ_BEST_SPEED;
    return SZ_SCES;
}

/* ref: sz/src/sz.c:75-95 */
int SZ_Init_Params(sz_params *params) {
    SZ_Init(NULL);
    if (params->losslessCompressor != GZIP_COMPRESSOR && params->losslessCompressor != ZSTD_COMPRESSOR) params->losslessCompressor = ZSTD_COMPRESSOR;
    if (params->max_quant_intervals > 0) params->maxRangeRadius = params->max_quant_intervals / 2;
    memcpy(confparams_cpr, params, sizeof(sz_params));
    if (params->quantization_intervals % 2 != 0) { printf("Error: quantization_intervals must be an even number!\n"); return SZ_NSCS; }
    return SZ_SCES;
}

/* ref: sz/src/sz.c:1296-1323 */
void SZ_Finalize(void) {
    if (confparams_dec != NULL) { free(confparams_dec); confparams_dec = NULL; }
    if (confparams_cpr != NULL) { free(confparams_cpr); confparams_cpr = NULL; }
    if (exe_params != NULL) { free(exe_params); exe_params = NULL; }
}

/* ref: sz/src/sz.c:97-115 */
int computeDimension(size_t r5, size_t r4, size_t r3, size_t r2, size_t r1) {
    if (r1 == 0) return 0;
    if (r2 == 0) return 1;
    if (r3 == 0) return 2;
    if (r4 == 0) return 3;
    if (r5 == 0) return 4;
    return 5;
}
/* ref: sz/src/sz.c:117-160 */
size_t computeDataLength(size_t r5, size_t r4, size_t r3, size_t r2, size_t r1) {
    if (r1 == 0) return 0;
    if (r2 == 0) return r1;
    if (r3 == 0) return r1 * r2;
    if (r4 == 0) return r1 * r2 * r3;
    if (r5 == 0) return r1 * r2 * r3 * r4;
    return r1 * r2 * r3 * r4 * r5;
}
/* ref: sz/src/sz.c:162-282: every dimension of extent 1 is squeezed out (order kept); c[0] = r1 ... c[4] = r5.
 * return 0 unchanged, 1 changed, 2 problematic */
int filterDimension(size_t r5, size_t r4, size_t r3, size_t r2, size_t r1, size_t *c) {
    size_t in[5] = {r1, r2, r3, r4, r5};
    int dim = computeDimension(r5, r4, r3, r2, r1), changed = 0, n = 0;
    for (int i = 0; i < 5; i++) c[i] = in[i];
    if (dim <= 1) return (dim == 1 && r1 < 1) ? 2 : 0;
    size_t out[5] = {0, 0, 0, 0, 0};
    for (int i = 0; i < dim; i++) { if (in[i] == 1) changed = 1; else out[n++] = in[i]; }
    if (changed) for (int i = 0; i < 5; i++) c[i] = out[i];
    return changed;
}

/* ref: sz/src/ByteToolkit.c:874-972 */
void convertSZParamsToBytes(sz_params *params, unsigned char *result) {
    unsigned char buf = (unsigned char)EXE()->optQuantMode;
    buf = (unsigned char)((buf << 1) | dataEndianType);
    buf = (unsigned char)((buf << 1) | sysEndianType);
    buf = (unsigned char)((buf << 2) | params->szMode);
    int tmp = 0;
    switch (params->gzipMode) { case 1: tmp = 0; break; case 0: tmp = 1; break; case 9: tmp = 2; break; }   /* Z_BEST_SPEED / Z_DEFAULT_STRATEGY / Z_BEST_COMPRESSION */
    buf = (unsigned char)((buf << 2) | tmp);
    result[0] = buf;
    put_be16(&result[1], (uint16_t)(int16_t)params->sampleDistance);
    short tmp2 = (short)(params->predThreshold * 10000);
    put_be16(&result[3], (uint16_t)tmp2);
    result[5] = (unsigned char)params->errorBoundMode;
    result[5] = (unsigned char)((result[5] << 4) | (params->dataType & 0x17));
    switch (params->errorBoundMode) {
    case ABS: put_be_float(&result[6], (float)params->absErrBound); memset(&result[10], 0, 4); break;
    case REL: memset(&result[6], 0, 4); put_be_float(&result[10], (float)params->relBoundRatio); break;
    case ABS_AND_REL:
    case ABS_OR_REL: put_be_float(&result[6], (float)params->absErrBound); put_be_float(&result[10], (float)params->relBoundRatio); break;
    case PSNR: put_be_float(&result[6], (float)params->psnr); memset(&result[9], 0, 4); break;
    case ABS_AND_PW_REL:
    case ABS_OR_PW_REL: put_be_float(&result[6], (float)params->absErrBound); put_be_float(&result[10], (float)params->pw_relBoundRatio); break;
    case REL_AND_PW_REL:
    case REL_OR_PW_REL: put_be_float(&result[6], (float)params->relBoundRatio); put_be_float(&result[10], (float)params->pw_relBoundRatio); break;
    case PW_REL: memset(&result[6], 0, 4); put_be_float(&result[10], (float)params->pw_relBoundRatio); break;
    }
    result[14] = (unsigned char)params->sol_ID;
    put_be32(&result[16], EXE()->optQuantMode == 1 ? params->max_quant_intervals : params->quantization_intervals);
    if (params->dataType == SZ_FLOAT) { put_be_float(&result[20], params->fmin); put_be_float(&result[24], params->fmax); }
    else { put_be_double(&result[20], params->dmin); put_be_double(&result[28], params->dmax); }
}

/* ref: sz/src/ByteToolkit.c:974-1063 */
void convertBytesToSZParams(unsigned char *bytes, sz_params *params) {
    unsigned char flag1 = bytes[0];
    EXE()->optQuantMode = (char)((flag1 & 0x40) >> 6);
    dataEndianType = (flag1 & 0x20) >> 5;
    params->szMode = (flag1 & 0x0c) >> 2;
    switch (flag1 & 0x03) { case 0: params->gzipMode = 1; break; case 1: params->gzipMode = 0; break; case 2: params->gzipMode = 9; break; }
    params->sampleDistance = (int16_t)get_be16(&bytes[1]);
    params->predThreshold = (float)(1.0 * (int16_t)get_be16(&bytes[3]) / 10000.0);
    params->dataType = bytes[5] & 0x07;
    params->errorBoundMode = (bytes[5] & 0xf0) >> 4;
    switch (params->errorBoundMode) {
    case ABS: params->absErrBound = get_be_float(&bytes[6]); break;
    case REL: params->relBoundRatio = get_be_float(&bytes[10]); break;
    case ABS_AND_REL:
    case ABS_OR_REL: params->absErrBound = get_be_float(&bytes[6]); params->relBoundRatio = get_be_float(&bytes[10]); break;
    case PSNR: params->psnr = get_be_float(&bytes[6]); break;
    case ABS_AND_PW_REL:
    case ABS_OR_PW_REL: params->absErrBound = get_be_float(&bytes[6]); params->pw_relBoundRatio = get_be_float(&bytes[10]); break;
    case REL_AND_PW_REL:
    case REL_OR_PW_REL: params->relBoundRatio = get_be_float(&bytes[6]); params->pw_relBoundRatio = get_be_float(&bytes[10]); break;
    case PW_REL: params->pw_relBoundRatio = get_be_float(&bytes[10]); break;
    }
    params->sol_ID = (int)bytes[14];
    if (EXE()->optQuantMode == 1) { params->max_quant_intervals = get_be32(&bytes[16]); params->quantization_intervals = 0; }
    else { params->max_quant_intervals = 0; params->quantization_intervals = get_be32(&bytes[16]); }
    if (params->dataType == SZ_FLOAT) { params->fmin = get_be_float(&bytes[20]); params->fmax = get_be_float(&bytes[24]); }
    else if (params->dataType == SZ_DOUBLE) { params->dmin = get_be_double(&bytes[20]); params->dmax = get_be_double(&bytes[28]); }
}

/* ref: sz/src/dataCompression.c:288-332 (getRealPrecision_double / _float; note min_f/max_f narrow BOTH operands) */
static double real_precision(int dataType, double range, int mode, double absb, double relb, int *status) {
    *status = SZ_SCES;
    if (mode == ABS) return absb;
    if (mode == REL) return relb * range;
    if (mode == ABS_AND_REL || mode == ABS_OR_REL) {
        if (dataType == SZ_FLOAT) {
            float a = (float)absb, b = (float)(relb * range);
            return (mode == ABS_AND_REL) ? (double)(a < b ? a : b) : (double)(a > b ? a : b);
        }
        double a = absb, b = relb * range;
        return (mode == ABS_AND_REL) ? (a < b ? a : b) : (a > b ? a : b);
    }
    *status = SZ_BERR;
    return 0;
}

static void lazy_init(void) {                          /* ref: sz/src/sz.c:297-309 */
    if (confparams_cpr == NULL) SZ_Init(NULL);
    else if (exe_params == NULL) exe_params = (sz_exedata *)calloc(1, sizeof(sz_exedata));
    if (exe_params->intvCapacity == 0) {
        exe_params->intvCapacity = (int)confparams_cpr->maxRangeRadius * 2;
        exe_params->intvRadius = (int)confparams_cpr->maxRangeRadius;
        exe_params->optQuantMode = 1;
    }
    if (exe_params->SZ_SIZE_TYPE == 0) exe_params->SZ_SIZE_TYPE = 8;
}

size_t szb200_compress_bound(int dataType, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1) {
    size_t r[5]; filterDimension(r5, r4, r3, r2, r1, r);
    size_t N = computeDataLength(r[4], r[3], r[2], r[1], r[0]);
    size_t ts = dataType == SZ_FLOAT ? 4 : 8, md = dataType == SZ_FLOAT ? MetaDataByteLength : MetaDataByteLength_double;
    szb200_job job; memset(&job, 0, sizeof(job));
    job.dtype = dataType;
    if (r[2] == 0) { job.dim = 2; job.n0 = 1; job.n1 = r[1] ? r[1] : 1; job.n2 = r[0] ? r[0] : 1; }
    else if (r[3] == 0) { job.dim = 3; job.n0 = r[2]; job.n1 = r[1]; job.n2 = r[0]; }
    else { job.dim = 3; job.n0 = r[3] * r[2]; job.n1 = r[1]; job.n2 = r[0]; }
    size_t body = szb200_body_capacity(&job), lossless = 4 + md + 8 + N * ts + 8;
    return 4 + md + (body > lossless ? body : lossless);
}

/* The common compress path.  Mirrors SZ_compress_args_float / _double (ref: sz_float.c:2811-3043, sz_double.c:2531-2740).
 * data: host pointer (data_on_device = 0) or device pointer.  Output either to a malloc'd host buffer (*host_out) or to
 * caller-provided device memory (dev_out, dev_cap). */
static int compress_core(int dataType, const void *data, int data_on_device, unsigned char **host_out, void *dev_out, size_t dev_cap,
                         size_t *outSize, int mode, double absb, double relb, const size_t *r) {
    const size_t r1 = r[0], r2 = r[1], r3 = r[2], r4 = r[3], r5 = r[4];
    const size_t ts = dataType == SZ_FLOAT ? 4 : 8;
    const size_t md = dataType == SZ_FLOAT ? MetaDataByteLength : MetaDataByteLength_double;
    CPR()->dataType = dataType;
    CPR()->errorBoundMode = mode;
    const size_t N = computeDataLength(r5, r4, r3, r2, r1);
    const int to_dev = (host_out == NULL);
    if (N <= MIN_NUM_OF_ELEMENTS) {                   /* SZ_skip_compress_*, ref: sz_float.c:37-43: raw bytes, no header */
        if (to_dev) { if (N * ts > dev_cap) return SZ_NSCS; if (N && szb200_copy(dev_out, data, N * ts, data_on_device ? 2 : 0)) return SZ_NSCS; }
        else { *host_out = (unsigned char *)malloc(N * ts + 1); if (N && (data_on_device ? szb200_copy(*host_out, data, N * ts, 1) : (memcpy(*host_out, data, N * ts), 0))) return SZ_NSCS; }
        *outSize = N * ts;
        return SZ_SCES;
    }
    if (mode != ABS && mode != REL && mode != ABS_AND_REL && mode != ABS_OR_REL && mode != PSNR && mode != NORM) {
        printf("Error: libSZ_b200 supports errorBoundMode ABS, REL, ABS_AND_REL, ABS_OR_REL, PSNR and NORM only (got %d).\n", mode);
        return SZ_BERR;
    }
    if (CPR()->szMode != SZ_BEST_SPEED) {
        printf("Error: libSZ_b200 implements szMode = SZ_BEST_SPEED only (the lossless back-end is out of scope); set szMode = SZ_BEST_SPEED.\n");
        return SZ_MERR;
    }
    if (r2 == 0 || CPR()->withRegression == SZ_NO_REGRESSION) {
        printf("Error: libSZ_b200 implements the SZ 2.1 blocked-regression path (2D/3D/4D, withLinearRegression = YES) only.\n");
        return SZ_DERR;
    }
    if (r5 != 0) { printf("Error: doesn't support 5 dimensions for now.\n"); return SZ_DERR; }
    if (CPR()->randomAccess) { printf("Error: randomAccess is not supported by libSZ_b200.\n"); return SZ_NSCS; }

    const int hprof = getenv("SZB200_HOST_PROFILE") != NULL;
    const double hp0 = hprof ? now_ms() : 0.0;
    const void *d_in = data;
    if (!data_on_device) { d_in = szb200_stage_input(data, N * ts); if (!d_in) { printf("Error: %s\n", szb200_last_error()); return SZ_NSCS; } }
    double vmin, vmax;
    if (szb200_range(dataType, d_in, N, &vmin, &vmax)) { printf("Error: %s\n", szb200_last_error()); return SZ_NSCS; }
    double range_d;
    if (dataType == SZ_FLOAT) {                        /* range and max are computed in T (ref: dataCompression.c:116, sz_float.c:2846) */
        float mn = (float)vmin, mx = (float)vmax, range = mx - mn, maxr = mn + range;
        CPR()->fmin = mn; CPR()->fmax = maxr; range_d = (double)range;
    } else {
        double range = vmax - vmin, maxr = vmin + range;
        CPR()->dmin = vmin; CPR()->dmax = maxr; range_d = range;
    }
    int status = SZ_SCES;
    double realPrecision;
    if (mode == PSNR) {                                /* ref: sz_float.c:2852-2856, conf.c:54-60: the stream is an ordinary ABS stream */
        CPR()->errorBoundMode = ABS;
        const double v1 = CPR()->psnr + 10 * log10(1 - 2.0 / 3.0 * (double)CPR()->predThreshold);
        const double v2 = v1 / (-20);
        const double v3 = pow(10, v2);
        realPrecision = range_d * v3;
    } else if (mode == NORM) {                         /* ref: sz_float.c:2858-2862, conf.c:62-65 */
        CPR()->errorBoundMode = ABS;
        realPrecision = sqrt(3.0 / N) * CPR()->normErr;
    } else realPrecision = real_precision(dataType, range_d, mode, absb, relb, &status);
    CPR()->absErrBound = realPrecision;

    unsigned char hdr[4 + MetaDataByteLength_double + 8 + 8];
    memset(hdr, 0, sizeof(hdr));
    hdr[0] = (unsigned char)versionNumber[0]; hdr[1] = (unsigned char)versionNumber[1]; hdr[2] = (unsigned char)versionNumber[2];

    if (range_d <= realPrecision) {                    /* constant container, ref: sz_float.c:2728-2753, TightDataPointStorageF.c:590-629 */
        unsigned char same = 0x01 | 0x40;
        if (CPR()->protectValueRange) same |= 0x04;
        hdr[3] = same;
        convertSZParamsToBytes(CPR(), hdr + 4);
        put_be64(hdr + 4 + md, (uint64_t)N);
        unsigned char first[8];
        if (data_on_device) { if (szb200_copy(first, data, ts, 1)) return SZ_NSCS; } else memcpy(first, data, ts);
        if (dataType == SZ_FLOAT) { float v; memcpy(&v, first, 4); put_be_float(hdr + 4 + md + 8, v); }
        else { double v; memcpy(&v, first, 8); put_be_double(hdr + 4 + md + 8, v); }
        size_t tot = 4 + md + 8 + ts;
        if (to_dev) { if (tot > dev_cap || szb200_copy(dev_out, hdr, tot, 0)) return SZ_NSCS; }
        else { *host_out = (unsigned char *)malloc(tot); memcpy(*host_out, hdr, tot); }
        *outSize = tot;
        return status;
    }

    szb200_job job; memset(&job, 0, sizeof(job));
    job.dtype = dataType;
    if (r3 == 0) { job.dim = 2; job.n0 = 1; job.n1 = r2; job.n2 = r1; }
    else if (r4 == 0) { job.dim = 3; job.n0 = r3; job.n1 = r2; job.n2 = r1; }
    else { job.dim = 3; job.n0 = r4 * r3; job.n1 = r2; job.n2 = r1; }          /* 4D flattened, ref: sz_float.c:3010 */
    job.eb = realPrecision;
    job.max_quant_intervals = CPR()->max_quant_intervals;
    job.quantization_intervals = EXE()->optQuantMode == 1 ? 0 : (unsigned int)EXE()->intvCapacity;
    job.sampleDistance = CPR()->sampleDistance;
    job.predThreshold = CPR()->predThreshold;

    const size_t lossless_total = 3 + md + 8 + 1 + ts * N;
    size_t cap = 4 + md + szb200_body_capacity(&job);
    if (cap < lossless_total + 16) cap = lossless_total + 16;
    unsigned char *out = NULL;       /* where the stream is assembled: the caller's device buffer, or the context's device staging buffer */
    if (to_dev) { if (dev_cap < lossless_total) { printf("Error: device stream buffer too small (need at least %zu bytes)\n", lossless_total); return SZ_NSCS; } cap = dev_cap; out = (unsigned char *)dev_out; }
    else { out = (unsigned char *)szb200_output_buffer(cap); if (!out) { printf("Error: %s\n", szb200_last_error()); return SZ_NSCS; } }
    size_t body = 0;
    szb200_stats st; memset(&st, 0, sizeof(st));
    const double hp1 = hprof ? now_ms() : 0.0;
    /* host destination: guess the size, allocate and pre-fault while the GPU works */
    unsigned char *early = NULL; size_t early_cap = 0; pthread_t pf_thread; int pf_started = 0; prefault_job pf;
    if (!to_dev && tl_cpr == NULL) {
        size_t guess = g_last_stream_bytes ? g_last_stream_bytes + g_last_stream_bytes / 4 : N * ts / 12;
        if (guess > lossless_total) guess = lossless_total;
        if (guess >= ((size_t)4 << 20)) {
            early = (unsigned char *)host_alloc_result(guess);
            if (early) { early_cap = guess; pf.p = early; pf.n = guess; pf_started = pthread_create(&pf_thread, NULL, prefault_thread, &pf) == 0; }
        }
    }
    int rc = szb200_compress_blocked(&job, d_in, out + 4 + md, 1, cap - 4 - md, &body, &st);
    if (pf_started) pthread_join(pf_thread, NULL);
    const double hp2 = hprof ? now_ms() : 0.0;
    /* A body that does not fit the caller's buffer is, in a buffer that holds the raw copy, exactly the case in which the reference
     * stores the original data instead (its trigger, sz_float.c:2979, is "stream at least as long as the raw copy"): not an error. */
    const int too_large = rc == SZB200_BODY_TOO_LARGE && 4 + md + body >= N * ts + 3 + md + 8 + 1;
    if (rc != SZ_SCES && !too_large) { printf("Error: %s\n", szb200_last_error()); free(early); return rc == SZB200_BODY_TOO_LARGE ? SZ_NSCS : rc; }
    if (!too_large && EXE()->optQuantMode == 1) { EXE()->intvCapacity = (int)st.intervals; EXE()->intvRadius = (int)st.intervals / 2; }   /* side effect of the reference (sz_float.c:6651) */
    size_t total = 4 + md + body;
    unsigned char same = 0x80 | 0x40;                   /* initRandomAccessBytes, ref: dataCompression.c:686-709 */
    if (CPR()->protectValueRange) same |= 0x04;
    const int lossless = total >= N * ts + 3 + md + 8 + 1;   /* SZ_compress_args_*_StoreOriData, ref: sz_float.c:526-559, trigger :2979 */
    if (lossless) { same = 80; total = lossless_total; }
    hdr[3] = same;
    convertSZParamsToBytes(CPR(), hdr + 4);
    unsigned char nb8[8]; put_be64(nb8, (uint64_t)N);
    if (to_dev) {
        if (lossless) {
            if (szb200_copy(out + 4 + md, nb8, 8, 0)) return SZ_NSCS;
            if (szb200_swap_copy(dataType, d_in, 1, out + 4 + md + 8, 1, N)) { printf("Error: %s\n", szb200_last_error()); return SZ_NSCS; }
        }
        if (szb200_copy(out, hdr, 4 + md, 0)) return SZ_NSCS;
    } else {
        /* the stream was assembled in device memory; the host result is allocated at its exact size and filled by ONE copy */
        unsigned char *h = NULL;
        if (early && total <= early_cap) { h = early; early = NULL; } else { free(early); early = NULL; h = (unsigned char *)host_alloc_result(total); }
        if (!h) return SZ_NSCS;
        g_last_stream_bytes = total;
        memcpy(h, hdr, 4 + md);
        if (lossless) {
            memcpy(h + 4 + md, nb8, 8);
            if (szb200_swap_copy(dataType, d_in, 1, h + 4 + md + 8, 0, N)) { printf("Error: %s\n", szb200_last_error()); free(h); return SZ_NSCS; }
        } else if (body && szb200_copy(h + 4 + md, out + 4 + md, body, 1)) { printf("Error: %s\n", szb200_last_error()); free(h); return SZ_NSCS; }
        *host_out = h;
    }
    *outSize = total;
    if (hprof) fprintf(stderr, "[host-prof] H2D + range %.2f ms | blocked compress call %.2f ms (device %.2f ms) | header + shrink %.2f ms\n", hp1 - hp0, hp2 - hp1, st.ms_total, now_ms() - hp2);
    return status;
}

/* ref: sz/src/sz.c:294-391 */
unsigned char *SZ_compress_args(int dataType, void *data, size_t *outSize, int errBoundMode, double absErrBound,
                                double relBoundRatio, double pwrBoundRatio, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1) {
    (void)pwrBoundRatio;
    pthread_mutex_lock(&g_api_mu);
    lazy_init();
    size_t r[5];
    filterDimension(r5, r4, r3, r2, r1, r);
    unsigned char *out = NULL;
    if (dataType == SZ_FLOAT || dataType == SZ_DOUBLE) {
        int rc = compress_core(dataType, data, 0, &out, NULL, 0, outSize, errBoundMode, absErrBound, relBoundRatio, r);
        if (rc != SZ_SCES) out = NULL;                 /* compress_core releases its buffer on every failure path */
    } else printf("Error: libSZ_b200 supports dataType SZ_FLOAT and SZ_DOUBLE only.\n");
    pthread_mutex_unlock(&g_api_mu);
    return out;
}

/* ref: sz/src/sz.c:393-401 */
int SZ_compress_args2(int dataType, void *data, unsigned char *compressed_bytes, size_t *outSize, int errBoundMode, double absErrBound,
                      double relBoundRatio, double pwrBoundRatio, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1) {
    unsigned char *bytes = SZ_compress_args(dataType, data, outSize, errBoundMode, absErrBound, relBoundRatio, pwrBoundRatio, r5, r4, r3, r2, r1);
    if (!bytes) return SZ_NSCS;
    memcpy(compressed_bytes, bytes, *outSize);
    free(bytes);
    return SZ_SCES;
}

/* ref: sz/src/sz.c:436-441 */
unsigned char *SZ_compress(int dataType, void *data, size_t *outSize, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1) {
    if (confparams_cpr == NULL) SZ_Init(NULL);
    return SZ_compress_args(dataType, data, outSize, confparams_cpr->errorBoundMode, confparams_cpr->absErrBound, confparams_cpr->relBoundRatio,
                            confparams_cpr->pw_relBoundRatio, r5, r4, r3, r2, r1);
}

int szb200_compress_args_device(int dataType, const void *d_data, void *d_stream, size_t stream_capacity, size_t *outSize,
                                int errBoundMode, double absErrBound, double relBoundRatio, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1) {
    if (dataType != SZ_FLOAT && dataType != SZ_DOUBLE) return SZ_NSCS;
    pthread_mutex_lock(&g_api_mu);
    lazy_init();
    size_t r[5];
    filterDimension(r5, r4, r3, r2, r1, r);
    int rc = compress_core(dataType, d_data, 1, NULL, d_stream, stream_capacity, outSize, errBoundMode, absErrBound, relBoundRatio, r);
    pthread_mutex_unlock(&g_api_mu);
    return rc;
}

/* ---- decompression ------------------------------------------------------------------------------------ */
static int looks_lossless_compressed(const unsigned char *b, size_t n) {   /* ref: sz/src/utility.c:156-172, callZlib.c:30-47 */
    if (n >= 4 && b[0] == 0x28 && b[1] == 0xB5 && b[2] == 0x2F && b[3] == 0xFD) return 1;   /* zstd frame magic */
    if (n >= 2) {
        if (b[0] == 104 && (b[1] == 5 || b[1] == 129 || b[1] == 222)) return 1;
        if (b[0] == 120 && (b[1] == 1 || b[1] == 94 || b[1] == 156 || b[1] == 218)) return 1;
    }
    return 0;
}

/* Mirrors SZ_decompress_args_float / _double (ref: szd_float.c:50-183) + new_TightDataPointStorageF_fromFlatBytes
 * (ref: TightDataPointStorageF.c:54-126).  bytes: host or device stream; output: device memory d_out (N elements). */
static int decompress_core(int dataType, const unsigned char *bytes, size_t len, int on_device, void *d_out, const size_t *r) {
    const size_t r1 = r[0], r2 = r[1], r3 = r[2], r4 = r[3], r5 = r[4];
    const size_t ts = dataType == SZ_FLOAT ? 4 : 8;
    const size_t md = dataType == SZ_FLOAT ? MetaDataByteLength : MetaDataByteLength_double;
    const size_t N = computeDataLength(r5, r4, r3, r2, r1);
    if (N <= MIN_NUM_OF_ELEMENTS && len == N * ts) {   /* raw container of SZ_skip_compress_* (the reference cannot decode its own; we can) */
        return N ? szb200_copy(d_out, bytes, len, on_device ? 2 : 0) : SZ_SCES;
    }
    if (len < 4 + md + 8) { printf("Error: compressed stream too short.\n"); return SZ_NSCS; }
    unsigned char head[4 + MetaDataByteLength_double + 8 + 8];
    size_t hl = 4 + md + 8 + ts; if (hl > len) hl = len;
    if (on_device) { if (szb200_copy(head, bytes, hl, 1)) return SZ_NSCS; } else memcpy(head, bytes, hl);
    if (looks_lossless_compressed(head, hl)) { printf("Error: stream was produced with a zstd/gzip szMode; libSZ_b200 decodes SZ_BEST_SPEED streams only.\n"); return SZ_MERR; }
    if (!(head[0] == versionNumber[0] && head[1] == versionNumber[1])) {   /* checkVersion2: major.minor must match (the reference exit(0)s here) */
        printf("Wrong version: \nCompressed-data version (%d.%d.%d)\nCurrent sz version: (%d.%d.%d)\n", head[0], head[1], head[2], versionNumber[0], versionNumber[1], versionNumber[2]);
        return SZ_NSCS;
    }
    const unsigned char same = head[3];
    EXE()->SZ_SIZE_TYPE = ((same & 0x40) >> 6) == 1 ? 8 : 4;
    if (EXE()->SZ_SIZE_TYPE != 8) { printf("Error: streams with a 4-byte size type are not supported.\n"); return SZ_NSCS; }
    DEC()->protectValueRange = (same & 0x04) >> 2;
    convertBytesToSZParams(head + 4, DEC());
    DEC()->szMode = SZ_BEST_SPEED;
    if (DEC()->dataType != dataType) { printf("Error: stream holds dataType %d, caller asked for %d.\n", DEC()->dataType, dataType); return SZ_NSCS; }
    if (DEC()->sol_ID != SZ) { printf("Error: unsupported sol_ID %d in stream.\n", DEC()->sol_ID); return SZ_NSCS; }
    const size_t Ns = (size_t)get_be64(head + 4 + md);
    if (Ns != N) { printf("Error: stream holds %zu elements, dimensions say %zu.\n", Ns, N); return SZ_DERR; }
    const unsigned char *body = bytes + 4 + md + 8;
    const size_t body_len = len - (4 + md + 8);
    int rc = SZ_SCES;
    if (same & 0x10) {                                 /* lossless container, ref: szd_float.c:106-119 */
        if (body_len < N * ts) { printf("Error: truncated stream.\n"); return SZ_NSCS; }
        rc = szb200_swap_copy(dataType, body, on_device, d_out, 1, N);
    } else if (same & 0x01) {                          /* constant container */
        if (body_len < ts) { printf("Error: truncated stream.\n"); return SZ_NSCS; }
        double v = dataType == SZ_FLOAT ? (double)get_be_float(head + 4 + md + 8) : get_be_double(head + 4 + md + 8);
        rc = szb200_fill(dataType, v, d_out, N);
    } else if (same & 0x80) {                          /* blocked-regression stream, ref: szd_float.c:126-141 */
        szb200_job job; memset(&job, 0, sizeof(job));
        job.dtype = dataType;
        if (r2 == 0) { printf("Error: 1D streams (SZ 1.4 container) are not supported by libSZ_b200.\n"); return SZ_DERR; }
        else if (r3 == 0) { job.dim = 2; job.n0 = 1; job.n1 = r2; job.n2 = r1; }
        else if (r4 == 0) { job.dim = 3; job.n0 = r3; job.n1 = r2; job.n2 = r1; }
        else if (r5 == 0) { job.dim = 3; job.n0 = r4 * r3; job.n1 = r2; job.n2 = r1; }
        else { printf("Error: currently support only at most 4 dimensions!\n"); return SZ_DERR; }
        rc = szb200_decompress_blocked(&job, body, body_len, on_device, d_out, NULL);
    } else { printf("Error: SZ 1.4 (non-regression) streams are not supported by libSZ_b200.\n"); return SZ_NSCS; }
    if (rc != SZ_SCES) { printf("Error: %s\n", szb200_last_error()); return rc; }
    if (DEC()->protectValueRange) {           /* ref: szd_float.c:161-176 */
        double mn = dataType == SZ_FLOAT ? (double)DEC()->fmin : DEC()->dmin;
        double mx = dataType == SZ_FLOAT ? (double)DEC()->fmax : DEC()->dmax;
        rc = szb200_clamp(dataType, d_out, N, mn, mx);
    }
    return rc;
}

static void reset_dec_state(void) {                   /* ref: sz/src/sz.c:497-510 */
    if (confparams_dec == NULL) confparams_dec = (sz_params *)malloc(sizeof(sz_params));
    memset(confparams_dec, 0, sizeof(sz_params));
    if (exe_params == NULL) exe_params = (sz_exedata *)malloc(sizeof(sz_exedata));
    memset(exe_params, 0, sizeof(sz_exedata));
    exe_params->SZ_SIZE_TYPE = 8;
    detect_endianness();
}

/* decode on the device, then ONE device-to-host copy into `dst` (NULL: a fresh malloc, returned through *made) */
static int decompress_to_host(int dataType, unsigned char *bytes, size_t byteLength, void *dst, void **made, const size_t r[5]) {
    const size_t ts = dataType == SZ_FLOAT ? 4 : 8;
    const size_t N = computeDataLength(r[4], r[3], r[2], r[1], r[0]);
    void *dbuf = szb200_output_buffer(N * ts + 64);
    if (!dbuf) return SZ_NSCS;
    /* a fresh result buffer is allocated first and helper threads touch its pages while the GPU decodes: the final copy is bound by
     * the host's page faults otherwise (512 MB: 37 ms of faults and copies against 10 ms of DMA) */
    enum { PF_THREADS = 4 };
    pthread_t pf_t[PF_THREADS]; prefault_job pf_j[PF_THREADS]; int pf_n = 0;
    if (!dst) {
        dst = host_alloc_result(N * ts); if (!dst) return SZ_NSCS; *made = dst;
        if (N * ts >= ((size_t)64 << 20)) {
            const size_t part = ((N * ts / PF_THREADS) + 4095) & ~(size_t)4095;
            for (int k = 0; k < PF_THREADS; k++) {
                const size_t off = (size_t)k * part; if (off >= N * ts) break;
                pf_j[pf_n].p = (unsigned char *)dst + off; pf_j[pf_n].n = off + part <= N * ts ? part : N * ts - off;
                if (pthread_create(&pf_t[pf_n], NULL, prefault_thread, &pf_j[pf_n]) == 0) pf_n++;
            }
        }
    }
    const int rc = decompress_core(dataType, bytes, byteLength, 0, dbuf, r);
    for (int k = 0; k < pf_n; k++) pthread_join(pf_t[k], NULL);
    if (rc != SZ_SCES || (N && szb200_copy(dst, dbuf, N * ts, 1) != SZ_SCES)) { if (made && *made) { free(*made); *made = NULL; } return SZ_NSCS; }
    return SZ_SCES;
}

/* ref: sz/src/sz.c:486-577 */
void *SZ_decompress(int dataType, unsigned char *bytes, size_t byteLength, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1) {
    if (dataType != SZ_FLOAT && dataType != SZ_DOUBLE) { printf("Error: data type cannot be the types other than SZ_FLOAT or SZ_DOUBLE\n"); return NULL; }
    pthread_mutex_lock(&g_api_mu);
    size_t r[5];
    filterDimension(r5, r4, r3, r2, r1, r);
    reset_dec_state();
    void *out = NULL;
    if (decompress_to_host(dataType, bytes, byteLength, NULL, &out, r) != SZ_SCES) out = NULL;
    pthread_mutex_unlock(&g_api_mu);
    return out;
}

/* ref: sz/src/sz.c:584-680.  The reference decompresses into a fresh array and copies; here the device result goes straight
 * into the caller's array (pinned memory makes this the fastest host-side call). */
size_t SZ_decompress_args(int dataType, unsigned char *bytes, size_t byteLength, void *decompressed_array, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1) {
    if (dataType != SZ_FLOAT && dataType != SZ_DOUBLE) { printf("Error: data type cannot be the types other than SZ_FLOAT or SZ_DOUBLE\n"); return 0; }
    if (!decompressed_array) return 0;
    pthread_mutex_lock(&g_api_mu);
    size_t r[5];
    filterDimension(r5, r4, r3, r2, r1, r);
    reset_dec_state();
    const size_t nbEle = computeDataLength(r[4], r[3], r[2], r[1], r[0]);
    const int rc = decompress_to_host(dataType, bytes, byteLength, decompressed_array, NULL, r);
    pthread_mutex_unlock(&g_api_mu);
    return rc == SZ_SCES ? nbEle : 0;
}

int szb200_decompress_device(int dataType, const void *d_stream, size_t byteLength, void *d_out, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1) {
    if (dataType != SZ_FLOAT && dataType != SZ_DOUBLE) return SZ_NSCS;
    pthread_mutex_lock(&g_api_mu);
    size_t r[5];
    filterDimension(r5, r4, r3, r2, r1, r);
    reset_dec_state();
    int rc = decompress_core(dataType, (const unsigned char *)d_stream, byteLength, 1, d_out, r);
    pthread_mutex_unlock(&g_api_mu);
    return rc;
}

/* ---- worker lanes for calls that run beside each other ---------------------------------------------------------------------
 * A lane is a context of the shim (streams + workspace) on the current device.  Lane 0 serves the ordinary calls, which the API
 * mutex serialises like the reference's global state demands; lanes 0 .. 7 serve the workers of a batch (which holds that mutex);
 * lanes 8 .. 15 are handed out here to the thread-safe calls (ref: sz/src/sz.c:1399, :1507) and to the slab workers: every such
 * call works on private copies of the parameter blocks and does not hold the API mutex while the GPU works, so the five threads
 * of example/test_threadsafe.c:40-58 really run beside each other. */
#define BATCH_MAX_LANES 8
#define POOL_FIRST_LANE 8
#define POOL_LANES 8
static pthread_mutex_t g_pool_mu = PTHREAD_MUTEX_INITIALIZER;
static pthread_cond_t g_pool_cv = PTHREAD_COND_INITIALIZER;
static int g_pool_busy[POOL_LANES];
static int pool_acquire(void) {
    pthread_mutex_lock(&g_pool_mu);
    for (;;) {
        for (int k = 0; k < POOL_LANES; k++) if (!g_pool_busy[k]) { g_pool_busy[k] = 1; pthread_mutex_unlock(&g_pool_mu); return POOL_FIRST_LANE + k; }
        pthread_cond_wait(&g_pool_cv, &g_pool_mu);
    }
}
static void pool_release(int lane) {
    pthread_mutex_lock(&g_pool_mu);
    g_pool_busy[lane - POOL_FIRST_LANE] = 0;
    pthread_cond_signal(&g_pool_cv);
    pthread_mutex_unlock(&g_pool_mu);
}

/* one compress / decompress on a pool lane with private parameter blocks (snapshots of the globals, taken under the API mutex) */
static int lane_compress(int dataType, const void *data, unsigned char **out, size_t *outSize, int mode, double absb, double relb, const size_t *r) {
    sz_params cpr; sz_exedata exe;
    pthread_mutex_lock(&g_api_mu);
    lazy_init();
    cpr = *confparams_cpr; exe = *exe_params;
    pthread_mutex_unlock(&g_api_mu);
    const int lane = pool_acquire();
    int rc = SZ_NSCS;
    if (szb200_set_lane(lane) == 0) {
        tl_cpr = &cpr; tl_exe = &exe;
        rc = compress_core(dataType, data, 0, out, NULL, 0, outSize, mode, absb, relb, r);
        tl_cpr = NULL; tl_exe = NULL;
    }
    szb200_set_lane(0);
    pool_release(lane);
    return rc;
}
static int lane_decompress(int dataType, unsigned char *bytes, size_t len, void *dst, void **made, const size_t *r) {
    sz_params dec; sz_exedata exe;
    memset(&dec, 0, sizeof(dec)); memset(&exe, 0, sizeof(exe)); exe.SZ_SIZE_TYPE = 8;           /* == reset_dec_state() */
    const int lane = pool_acquire();
    int rc = SZ_NSCS;
    if (szb200_set_lane(lane) == 0) {
        tl_dec = &dec; tl_exe = &exe;
        rc = decompress_to_host(dataType, bytes, len, dst, made, r);
        tl_dec = NULL; tl_exe = NULL;
    }
    szb200_set_lane(0);
    pool_release(lane);
    return rc;
}

/* ---- one array over several GPUs: slabs (include/szb200.h) ------------------------------------------------------------------
 * The slowest dimension is cut into n_slabs contiguous pieces (sizes floor or floor + 1, the larger ones first); slab s is
 * compressed on device s mod n_devices by a worker thread on a pool lane, as an independent SZ stream that equals the reference
 * called on that slab (the reference's own pattern: example/test_threadsafe.c:40-58).  Container (ours; the reference has none):
 *   [n_slabs : u32 LE][size_s : u64 LE x n_slabs][stream_0 ... stream_{n_slabs-1}]      -- the framing of sz_b200/slabs.py */
typedef struct slab_shared {
    int dataType, decompress, mode, n_slabs, n_devices;
    double absb, relb;
    size_t row_elems, n0;                       /* elements per row of the slowest dimension; its extent */
    size_t rlow[2];                             /* r1, r2 of every slab (a 4D array is cut as (r4 * r3, r2, r1), which the reference makes of it anyway) */
    const unsigned char *in;                    /* compress: the array; decompress: the container */
    unsigned char *out;                         /* decompress: the array */
    unsigned char **streams; size_t *sizes, *offs;
    int next, failed;
    pthread_mutex_t mu;
} slab_shared;
static void slab_range(const slab_shared *sh, int s, size_t *lo, size_t *hi) {
    const size_t base = sh->n0 / (size_t)sh->n_slabs, extra = sh->n0 % (size_t)sh->n_slabs;
    *lo = (size_t)s * base + ((size_t)s < extra ? (size_t)s : extra);
    *hi = *lo + base + ((size_t)s < extra ? 1 : 0);
}
static void *slab_thread(void *arg) {
    slab_shared *sh = (slab_shared *)arg;
    const size_t ts = sh->dataType == SZ_FLOAT ? 4 : 8;
    for (;;) {
        pthread_mutex_lock(&sh->mu);
        const int s = sh->next < sh->n_slabs ? sh->next++ : -1;
        pthread_mutex_unlock(&sh->mu);
        if (s < 0) break;
        size_t lo, hi; slab_range(sh, s, &lo, &hi);
        size_t r[5] = {sh->rlow[0], sh->rlow[1] ? sh->rlow[1] : hi - lo, sh->rlow[1] ? hi - lo : 0, 0, 0};   /* (r1, r2, rows) or, for a 2D array, (r1, rows) */
        size_t rf[5]; filterDimension(r[4], r[3], r[2], r[1], r[0], rf);
        int rc = SZ_NSCS;
        if (szb200_set_device(s % sh->n_devices) == 0) {
            if (!sh->decompress) rc = lane_compress(sh->dataType, sh->in + lo * sh->row_elems * ts, &sh->streams[s], &sh->sizes[s], sh->mode, sh->absb, sh->relb, rf);
            else rc = lane_decompress(sh->dataType, (unsigned char *)sh->in + sh->offs[s], sh->sizes[s], sh->out + lo * sh->row_elems * ts, NULL, rf);
        }
        if (rc != SZ_SCES) { pthread_mutex_lock(&sh->mu); sh->failed = 1; pthread_mutex_unlock(&sh->mu); }
    }
    return NULL;
}
static int slab_setup(slab_shared *sh, int dataType, int n_slabs, int n_devices, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1) {
    if (dataType != SZ_FLOAT && dataType != SZ_DOUBLE) return SZ_NSCS;
    size_t r[5]; filterDimension(r5, r4, r3, r2, r1, r);
    if (r[4] != 0 || r[1] == 0) { printf("Error: slabs need a 2D, 3D or 4D array.\n"); return SZ_DERR; }
    memset(sh, 0, sizeof(*sh));
    sh->dataType = dataType;
    if (r[2] == 0) { sh->n0 = r[1]; sh->rlow[0] = r[0]; sh->rlow[1] = 0; sh->row_elems = r[0]; }                       /* 2D: cut r2 */
    else { sh->n0 = r[3] ? r[3] * r[2] : r[2]; sh->rlow[0] = r[0]; sh->rlow[1] = r[1]; sh->row_elems = r[0] * r[1]; }  /* 3D / 4D */
    const int ndev = szb200_device_count();
    if (ndev <= 0) { printf("Error: no CUDA device.\n"); return SZ_NSCS; }
    sh->n_devices = (n_devices <= 0 || n_devices > ndev) ? ndev : n_devices;
    sh->n_slabs = n_slabs <= 0 ? sh->n_devices : n_slabs;
    if ((size_t)sh->n_slabs > sh->n0 / 12) sh->n_slabs = sh->n0 / 12 ? (int)(sh->n0 / 12) : 1;      /* at least two blocks of rows per slab */
    pthread_mutex_init(&sh->mu, NULL);
    return SZ_SCES;
}
static int slab_run(slab_shared *sh) {
    const int device0 = szb200_current_device();
    int workers = sh->n_slabs < POOL_LANES ? sh->n_slabs : POOL_LANES;
    pthread_t th[POOL_LANES]; int started = 0;
    for (int k = 0; k < workers; k++) { if (pthread_create(&th[k], NULL, slab_thread, sh) != 0) break; started++; }
    for (int k = 0; k < started; k++) pthread_join(th[k], NULL);
    pthread_mutex_destroy(&sh->mu);
    if (device0 >= 0) szb200_set_device(device0);
    return (started > 0 && !sh->failed && sh->next >= sh->n_slabs) ? SZ_SCES : SZ_NSCS;
}

int szb200_compress_slabs(int dataType, const void *data, int n_slabs, int n_devices, unsigned char **container, size_t *containerSize,
                          int errBoundMode, double absErrBound, double relBoundRatio, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1) {
    slab_shared sh;
    if (!data || !container || !containerSize) return SZ_NSCS;
    int rc = slab_setup(&sh, dataType, n_slabs, n_devices, r5, r4, r3, r2, r1);
    if (rc != SZ_SCES) return rc;
    sh.mode = errBoundMode; sh.absb = absErrBound; sh.relb = relBoundRatio; sh.in = (const unsigned char *)data;
    sh.streams = (unsigned char **)calloc((size_t)sh.n_slabs, sizeof(unsigned char *));
    sh.sizes = (size_t *)calloc((size_t)sh.n_slabs, sizeof(size_t));
    rc = (sh.streams && sh.sizes) ? slab_run(&sh) : SZ_NSCS;
    if (rc == SZ_SCES) {
        size_t total = 4 + 8 * (size_t)sh.n_slabs;
        for (int s = 0; s < sh.n_slabs; s++) total += sh.sizes[s];
        unsigned char *c = (unsigned char *)host_alloc_result(total);
        if (!c) rc = SZ_NSCS;
        else {
            const uint32_t g = (uint32_t)sh.n_slabs;
            memcpy(c, &g, 4);                                     /* little endian, like sz_b200/slabs.py */
            size_t off = 4 + 8 * (size_t)sh.n_slabs;
            for (int s = 0; s < sh.n_slabs; s++) { const uint64_t z = sh.sizes[s]; memcpy(c + 4 + 8 * (size_t)s, &z, 8); memcpy(c + off, sh.streams[s], sh.sizes[s]); off += sh.sizes[s]; }
            *container = c; *containerSize = total;
        }
    }
    if (sh.streams) for (int s = 0; s < sh.n_slabs; s++) free(sh.streams[s]);
    free(sh.streams); free(sh.sizes);
    return rc;
}

int szb200_decompress_slabs(int dataType, const unsigned char *container, size_t containerSize, int n_devices, void *out,
                            size_t r5, size_t r4, size_t r3, size_t r2, size_t r1) {
    slab_shared sh;
    if (!container || !out || containerSize < 4) return SZ_NSCS;
    uint32_t g; memcpy(&g, container, 4);
    if (g == 0 || g > 65536 || containerSize < 4 + 8 * (size_t)g) { printf("Error: not a slab container.\n"); return SZ_NSCS; }
    int rc = slab_setup(&sh, dataType, (int)g, n_devices, r5, r4, r3, r2, r1);
    if (rc != SZ_SCES) return rc;
    if (sh.n_slabs != (int)g) { printf("Error: the container holds %u slabs, which does not fit these dimensions.\n", g); pthread_mutex_destroy(&sh.mu); return SZ_DERR; }
    sh.decompress = 1; sh.in = container; sh.out = (unsigned char *)out;
    sh.sizes = (size_t *)calloc(g, sizeof(size_t)); sh.offs = (size_t *)calloc(g, sizeof(size_t));
    rc = (sh.sizes && sh.offs) ? SZ_SCES : SZ_NSCS;
    size_t off = 4 + 8 * (size_t)g;
    for (uint32_t s = 0; rc == SZ_SCES && s < g; s++) {
        uint64_t z; memcpy(&z, container + 4 + 8 * (size_t)s, 8);
        if (z > containerSize || off + z > containerSize) { printf("Error: truncated slab container.\n"); rc = SZ_NSCS; break; }
        sh.sizes[s] = (size_t)z; sh.offs[s] = off; off += (size_t)z;
    }
    if (rc == SZ_SCES) rc = slab_run(&sh); else pthread_mutex_destroy(&sh.mu);
    free(sh.sizes); free(sh.offs);
    return rc;
}

/* ---- batches (include/szb200.h) ------------------------------------------------------------------------------------------- */
typedef struct batch_shared {
    int dataType, decompress, mode, device;
    double absb, relb;
    size_t r[5];
    szb200_batch_item *items; int n_items;
    int next;                                   /* guarded by mu */
    pthread_mutex_t mu;
    sz_params cpr0; sz_exedata exe0;            /* snapshot of the caller's configuration */
} batch_shared;
typedef struct batch_worker { batch_shared *sh; int lane; } batch_worker;

static void *batch_thread(void *arg) {
    batch_worker *w = (batch_worker *)arg;
    batch_shared *sh = w->sh;
    sz_params cpr, dec; sz_exedata exe;
    if (szb200_set_device(sh->device) != 0 || szb200_set_lane(w->lane) != 0 || szb200_set_device(sh->device) != 0) return NULL;
    tl_cpr = &cpr; tl_dec = &dec; tl_exe = &exe;
    for (;;) {
        pthread_mutex_lock(&sh->mu);
        const int i = sh->next < sh->n_items ? sh->next++ : -1;
        pthread_mutex_unlock(&sh->mu);
        if (i < 0) break;
        szb200_batch_item *it = &sh->items[i];
        if (!sh->decompress) {
            cpr = sh->cpr0; exe = sh->exe0;     /* every item starts from the caller's configuration, like a fresh call */
            size_t n = 0;
            it->rc = compress_core(sh->dataType, it->d_data, 1, NULL, it->d_stream, it->stream_capacity, &n, sh->mode, sh->absb, sh->relb, sh->r);
            it->out_size = n;
        } else {
            memset(&dec, 0, sizeof(dec)); memset(&exe, 0, sizeof(exe)); exe.SZ_SIZE_TYPE = 8;      /* == reset_dec_state() */
            it->rc = decompress_core(sh->dataType, (const unsigned char *)it->d_stream, it->stream_capacity, 1, it->d_out, sh->r);
        }
    }
    tl_cpr = tl_dec = NULL; tl_exe = NULL;
    szb200_set_lane(0);
    return NULL;
}

static int run_batch(batch_shared *sh, int lanes) {
    if (lanes < 1) lanes = 1;
    if (lanes > BATCH_MAX_LANES) lanes = BATCH_MAX_LANES;     /* lanes 0 .. 7; 8 .. 15 belong to the thread-safe calls and the slab workers */
    if (lanes > sh->n_items) lanes = sh->n_items;
    pthread_t th[16]; batch_worker wk[16];
    pthread_mutex_init(&sh->mu, NULL);
    sh->next = 0;
    int started = 0;
    for (int k = 0; k < lanes; k++) { wk[k].sh = sh; wk[k].lane = k; if (pthread_create(&th[k], NULL, batch_thread, &wk[k]) != 0) break; started++; }
    for (int k = 0; k < started; k++) pthread_join(th[k], NULL);
    pthread_mutex_destroy(&sh->mu);
    if (started == 0) return SZ_NSCS;
    int rc = SZ_SCES;
    for (int i = 0; i < sh->n_items; i++) if (sh->items[i].rc != SZ_SCES) rc = sh->items[i].rc;
    return sh->next >= sh->n_items ? rc : SZ_NSCS;
}

int szb200_compress_batch_device(int dataType, szb200_batch_item *items, int n_items, int lanes, int errBoundMode, double absErrBound,
                                 double relBoundRatio, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1) {
    if ((dataType != SZ_FLOAT && dataType != SZ_DOUBLE) || !items || n_items <= 0) return SZ_NSCS;
    batch_shared sh; memset(&sh, 0, sizeof(sh));
    pthread_mutex_lock(&g_api_mu);
    lazy_init();
    sh.cpr0 = *confparams_cpr; sh.exe0 = *exe_params;
    sh.dataType = dataType; sh.decompress = 0; sh.mode = errBoundMode; sh.absb = absErrBound; sh.relb = relBoundRatio;
    sh.device = szb200_current_device();
    filterDimension(r5, r4, r3, r2, r1, sh.r);
    sh.items = items; sh.n_items = n_items;
    for (int i = 0; i < n_items; i++) { items[i].rc = SZ_NSCS; items[i].out_size = 0; }
    const int rc = sh.device < 0 ? SZ_NSCS : run_batch(&sh, lanes);
    pthread_mutex_unlock(&g_api_mu);
    return rc;
}

int szb200_decompress_batch_device(int dataType, szb200_batch_item *items, int n_items, int lanes,
                                   size_t r5, size_t r4, size_t r3, size_t r2, size_t r1) {
    if ((dataType != SZ_FLOAT && dataType != SZ_DOUBLE) || !items || n_items <= 0) return SZ_NSCS;
    batch_shared sh; memset(&sh, 0, sizeof(sh));
    pthread_mutex_lock(&g_api_mu);
    sh.dataType = dataType; sh.decompress = 1;
    sh.device = szb200_current_device();
    filterDimension(r5, r4, r3, r2, r1, sh.r);
    sh.items = items; sh.n_items = n_items;
    for (int i = 0; i < n_items; i++) items[i].rc = SZ_NSCS;
    const int rc = sh.device < 0 ? SZ_NSCS : run_batch(&sh, lanes);
    pthread_mutex_unlock(&g_api_mu);
    return rc;
}

/* ref: sz/src/sz.c:683-766 */
sz_metadata *SZ_getMetadata(unsigned char *bytes) {
    int index = 0;
    int versions[3];
    for (int i = 0; i < 3; i++) versions[i] = bytes[index++];
    unsigned char same = bytes[index++];
    int isConstant = same & 0x01, isLossless = (same & 0x10) >> 4, isRegressionBased = (same >> 7) & 0x01;
    if (exe_params == NULL) exe_params = (sz_exedata *)calloc(1, sizeof(sz_exedata));
    exe_params->SZ_SIZE_TYPE = ((same & 0x40) >> 6) == 1 ? 8 : 4;
    if (confparams_dec == NULL) confparams_dec = (sz_params *)calloc(1, sizeof(sz_params));
    convertBytesToSZParams(&bytes[index], confparams_dec);
    if (confparams_dec->dataType == SZ_FLOAT) index += MetaDataByteLength;
    else if (confparams_dec->dataType == SZ_DOUBLE) index += MetaDataByteLength_double;
    size_t n = exe_params->SZ_SIZE_TYPE == 8 ? (size_t)get_be64(&bytes[index]) : (size_t)get_be32(&bytes[index]);
    index += (int)exe_params->SZ_SIZE_TYPE;
    sz_metadata *m = (sz_metadata *)malloc(sizeof(sz_metadata));
    m->versionNumber[0] = versions[0]; m->versionNumber[1] = versions[1]; m->versionNumber[2] = versions[2];
    m->isConstant = isConstant; m->isLossless = isLossless; m->sizeType = (int)exe_params->SZ_SIZE_TYPE; m->dataSeriesLength = n;
    m->conf_params = confparams_dec;
    int bins = 0;
    /* the reference reads at raBytes + 4 + sizeof(double) for BOTH types (sz.c:741-745), which is only right for double
     * streams (SURVEY.md 8b "quirks"); here the precision field's real width is used. */
    if (!isConstant && !isLossless && isRegressionBased) bins = (int)get_be32(&bytes[index] + 4 + (confparams_dec->dataType == SZ_FLOAT ? 4 : 8));
    m->defactoNBBins = bins;
    return m;
}

/* ref: sz/src/sz.c:768-923 (same fields, same order) */
void SZ_printMetadata(sz_metadata *metadata) {
    sz_params *p = metadata->conf_params;
    printf("=================SZ Compression Meta Data=================\n");
    printf("Version:                        \t %d.%d.%d\n", metadata->versionNumber[0], metadata->versionNumber[1], metadata->versionNumber[2]);
    printf("Constant data?:                 \t %s\n", metadata->isConstant == 1 ? "YES" : "NO");
    printf("Lossless?:                      \t %s\n", metadata->isLossless == 1 ? "YES" : "NO");
    printf("Size type (size of # elements): \t %d bytes\n", metadata->sizeType);
    printf("Num of elements:                \t %zu\n", metadata->dataSeriesLength);
    printf("compressor Name: \t\t\t %s\n", p->sol_ID == SZ ? "SZ" : (p->sol_ID == SZ_Transpose ? "SZ_Transpose" : "Other compressor"));
    if (p->dataType == SZ_FLOAT) { printf("Data type:                      \t FLOAT\nmin value of raw data:          \t %f\nmax value of raw data:          \t %f\n", p->fmin, p->fmax); }
    else if (p->dataType == SZ_DOUBLE) { printf("Data type:                      \t DOUBLE\nmin value of raw data:          \t %f\nmax value of raw data:          \t %f\n", p->dmin, p->dmax); }
    if (exe_params->optQuantMode == 1) { printf("quantization_intervals:         \t 0\nmax_quant_intervals:            \t %u\n", p->max_quant_intervals); printf("actual used # intervals:        \t %d\n", metadata->defactoNBBins); }
    else printf("quantization_intervals:         \t %u\n", p->quantization_intervals);
    printf("dataEndianType (prior raw data):\t %s\n", dataEndianType == BIG_ENDIAN_DATA ? "BIG_ENDIAN" : "LITTLE_ENDIAN");
    printf("sysEndianType (at compression): \t %s\n", sysEndianType == 1 ? "BIG_ENDIAN" : "LITTLE_ENDIAN");
    printf("sampleDistance:                 \t %d\n", p->sampleDistance);
    printf("predThreshold:                  \t %f\n", p->predThreshold);
    printf("szMode:                         \t %s\n", p->szMode == SZ_BEST_SPEED ? "SZ_BEST_SPEED (without Gzip)" : "SZ_BEST_COMPRESSION / SZ_DEFAULT_COMPRESSION");
    switch (p->errorBoundMode) {
    case ABS: printf("errBoundMode:                   \t ABS\nabsErrBound:                    \t %f\n", p->absErrBound); break;
    case REL: printf("errBoundMode:                   \t REL (based on value_range extent)\nrelBoundRatio:                  \t %f\n", p->relBoundRatio); break;
    case ABS_AND_REL: printf("errBoundMode:                   \t ABS_AND_REL\nabsErrBound:                    \t %f\nrelBoundRatio:                  \t %f\n", p->absErrBound, p->relBoundRatio); break;
    case ABS_OR_REL: printf("errBoundMode:                   \t ABS_OR_REL\nabsErrBound:                    \t %f\nrelBoundRatio:                  \t %f\n", p->absErrBound, p->relBoundRatio); break;
    default: printf("errBoundMode:                   \t %d\n", p->errorBoundMode); break;
    }
}

/* ---- customize API, ref: sz/src/sz.c:1337-1510 (names "SZ", "SZ2.0", "SZ2.1") ---------------------------------- */
static int is_sz2_name(const char *n) { return strcmp(n, "SZ2.0") == 0 || strcmp(n, "SZ2.1") == 0 || strcmp(n, "SZ") == 0; }

unsigned char *SZ_compress_customize(const char *cmprName, void *userPara, int dataType, void *data, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1, size_t *outSize, int *status) {
    if (!is_sz2_name(cmprName)) { *status = SZ_NSCS; return NULL; }
    if (userPara == NULL && confparams_cpr == NULL) SZ_Init(NULL);
    else if (userPara != NULL) SZ_Init_Params((sz_params *)userPara);
    unsigned char *res = SZ_compress(dataType, data, outSize, r5, r4, r3, r2, r1);
    *status = res ? SZ_SCES : SZ_NSCS;
    return res;
}

unsigned char *SZ_compress_customize_threadsafe(const char *cmprName, void *userPara, int dataType, void *data, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1, size_t *outSize, int *status) {
    if (!is_sz2_name(cmprName) || userPara == NULL || (dataType != SZ_FLOAT && dataType != SZ_DOUBLE)) { *status = SZ_NSCS; return NULL; }
    sz_params *para = (sz_params *)userPara;
    /* ref: sz/src/sz.c:1399-1417 -- mode and bounds from the caller's block, everything else from the global configuration; the
     * call runs on a worker lane of its own with private parameter blocks (see "worker lanes" above), beside other such calls */
    size_t r[5];
    filterDimension(r5, r4, r3, r2, r1, r);
    unsigned char *res = NULL;
    const int rc = lane_compress(dataType, data, &res, outSize, para->errorBoundMode, para->absErrBound, para->relBoundRatio, r);
    if (rc != SZ_SCES) res = NULL;
    *status = res ? SZ_SCES : SZ_NSCS;
    return res;
}

void *SZ_decompress_customize(const char *cmprName, void *userPara, int dataType, unsigned char *bytes, size_t byteLength, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1, int *status) {
    (void)userPara;
    if (!(is_sz2_name(cmprName) || strcmp(cmprName, "SZ1.4") == 0)) { *status = SZ_NSCS; return NULL; }
    void *res = SZ_decompress(dataType, bytes, byteLength, r5, r4, r3, r2, r1);
    *status = res ? SZ_SCES : SZ_NSCS;
    return res;
}

void *SZ_decompress_customize_threadsafe(const char *cmprName, void *userPara, int dataType, unsigned char *bytes, size_t byteLength, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1, int *status) {
    (void)userPara;
    if (!(is_sz2_name(cmprName)) || (dataType != SZ_FLOAT && dataType != SZ_DOUBLE)) { *status = SZ_NSCS; return NULL; }
    size_t r[5];
    filterDimension(r5, r4, r3, r2, r1, r);
    void *out = NULL;
    if (lane_decompress(dataType, bytes, byteLength, NULL, &out, r) != SZ_SCES) out = NULL;
    *status = out ? SZ_SCES : SZ_NSCS;
    return out;
}
