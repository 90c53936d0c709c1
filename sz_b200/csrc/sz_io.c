/*
 * sz_io.c -- the handful of file helpers of the reference's sz/include/rw.h that its example programs call
 * (example/testfloat_compress.c, testfloat_decompress.c, testdouble_compress.c, testdouble_decompress.c), so that those
 * programs relink against libSZ_b200.so unmodified (recipe: oracle/Makefile target `dropin`).
 * Plain stdio on whole files; values are read / written in the byte order named by dataEndianType (ref: sz/src/rw.c:425-470,
 * :745-790, :864-905).  Nothing here is on the compression path.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>
#include "sz.h"

static void swap_elems(unsigned char *p, size_t n, size_t width) {
    for (size_t i = 0; i < n; i++, p += width)
        for (size_t a = 0, b = width - 1; a < b; a++, b--) { unsigned char t = p[a]; p[a] = p[b]; p[b] = t; }
}

/* ref: sz/src/rw.c:96-133 */
unsigned char *readByteData(char *srcFilePath, size_t *byteLength, int *status) {
    FILE *f = fopen(srcFilePath, "rb");
    if (!f) { printf("Failed to open input file. 1\n"); *status = SZ_FERR; return NULL; }
    fseek(f, 0, SEEK_END);
    const long len = ftell(f);
    fseek(f, 0, SEEK_SET);
    if (len < 0) { fclose(f); *status = SZ_FERR; return NULL; }
    unsigned char *buf = (unsigned char *)malloc((size_t)len ? (size_t)len : 1);
    if (!buf || fread(buf, 1, (size_t)len, f) != (size_t)len) { printf("Error: input file is wrong!\n"); free(buf); fclose(f); *status = SZ_FERR; return NULL; }
    fclose(f);
    *byteLength = (size_t)len;
    *status = SZ_SCES;
    return buf;
}

/* ref: sz/src/rw.c:730-743 */
void writeByteData(unsigned char *bytes, size_t byteLength, char *tgtFilePath, int *status) {
    FILE *f = fopen(tgtFilePath, "wb");
    if (!f) { printf("Failed to open input file. 3\n"); *status = SZ_FERR; return; }
    const size_t w = fwrite(bytes, 1, byteLength, f);
    fclose(f);
    *status = (w == byteLength) ? SZ_SCES : SZ_FERR;
}

static void *read_elems(char *path, size_t width, size_t *nbEle, int *status) {
    size_t len = 0;
    unsigned char *b = readByteData(path, &len, status);
    if (!b) return NULL;
    *nbEle = len / width;
    if (dataEndianType != sysEndianType) swap_elems(b, *nbEle, width);
    return b;
}
/* ref: sz/src/rw.c:425-461 */
float *readFloatData(char *srcFilePath, size_t *nbEle, int *status) { return (float *)read_elems(srcFilePath, 4, nbEle, status); }
/* ref: sz/src/rw.c:135-172 */
double *readDoubleData(char *srcFilePath, size_t *nbEle, int *status) { return (double *)read_elems(srcFilePath, 8, nbEle, status); }

/* ref: sz/src/rw.c:864-905: the values are written in the host's byte order */
void writeFloatData_inBytes(float *data, size_t nbEle, char *tgtFilePath, int *status) { writeByteData((unsigned char *)data, nbEle * sizeof(float), tgtFilePath, status); }
void writeDoubleData_inBytes(double *data, size_t nbEle, char *tgtFilePath, int *status) { writeByteData((unsigned char *)data, nbEle * sizeof(double), tgtFilePath, status); }

/* ---- the further helpers the reference's command-line tool (example/sz.c) calls, so that `sz` too relinks unmodified ---- */

/* ref: sz/src/rw.c:22-31 */
int checkFileExistance(char *filePath) { return access(filePath, F_OK) != -1 ? 1 : 0; }

/* ref: sz/src/rw.c:989-1009: one string per line */
void writeStrings(int nbStr, char *str[], char *tgtFilePath, int *status) {
    FILE *f = fopen(tgtFilePath, "wb");
    if (!f) { printf("Failed to open input file. 3\n"); *status = SZ_FERR; return; }
    for (int i = 0; i < nbStr; i++) { fputs(str[i], f); fputc('\n', f); }
    fclose(f);
    *status = SZ_SCES;
}

/* ref: sz/src/rw.c:796-840: text output, one value per line ("%.30G" for float, "%.20G" for double) */
void writeFloatData(float *data, size_t nbEle, char *tgtFilePath, int *status) {
    FILE *f = fopen(tgtFilePath, "wb");
    if (!f) { printf("Failed to open input file. 3\n"); *status = SZ_FERR; return; }
    for (size_t i = 0; i < nbEle; i++) fprintf(f, "%.30G\n", data[i]);
    fclose(f);
    *status = SZ_SCES;
}
void writeDoubleData(double *data, size_t nbEle, char *tgtFilePath, int *status) {
    FILE *f = fopen(tgtFilePath, "wb");
    if (!f) { printf("Failed to open input file. 3\n"); *status = SZ_FERR; return; }
    for (size_t i = 0; i < nbEle; i++) fprintf(f, "%.20G\n", data[i]);
    fclose(f);
    *status = SZ_SCES;
}

/* ref: sz/src/utility.c:236-330.  The reference's index loops are a plain transpose of an M x F matrix, F being the extent of the
 * fastest axis of the input (r1 in 2D, r3 in 3D, r4 in 4D) and M the rest; 1D is a copy. */
void *detransposeData(void *data, int dataType, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1) {
    const size_t len = computeDataLength(r5, r4, r3, r2, r1);
    const int dim = computeDimension(r5, r4, r3, r2, r1);
    const size_t ts = dataType == SZ_FLOAT ? 4 : 8;
    if (dataType != SZ_FLOAT && dataType != SZ_DOUBLE) return NULL;
    unsigned char *out = (unsigned char *)malloc(ts * (len ? len : 1));
    const unsigned char *in = (const unsigned char *)data;
    const size_t F = dim == 2 ? r1 : dim == 3 ? r3 : dim == 4 ? r4 : 1;
    if (!out) return NULL;
    if (dim < 2 || dim > 4 || F == 0) { memcpy(out, in, ts * len); return out; }
    const size_t M = len / F;
    for (size_t m = 0; m < M; m++)
        for (size_t f = 0; f < F; f++) memcpy(out + ts * (f * M + m), in + ts * (m * F + f), ts);
    return out;
}

/* ref: sz/src/utility.c:156-172, callZlib.c:30-47: does a file start like a zstd frame or a zlib stream?  (Streams written with
 * szMode = SZ_BEST_SPEED -- the only mode of this library -- start with the version bytes and are neither.) */
int is_lossless_compressed_data(unsigned char *compressedBytes, size_t cmpSize) {
    if (cmpSize >= 4 && compressedBytes[0] == 0x28 && compressedBytes[1] == 0xB5 && compressedBytes[2] == 0x2F && compressedBytes[3] == 0xFD) return ZSTD_COMPRESSOR;
    if (cmpSize >= 2) {
        const unsigned char a = compressedBytes[0], b = compressedBytes[1];
        if ((a == 104 && (b == 5 || b == 129 || b == 222)) || (a == 120 && (b == 1 || b == 94 || b == 156 || b == 218))) return GZIP_COMPRESSOR;
    }
    return -1;
}

/* ref: sz/src/utility.c:203-234.  The lossless back-end is outside this library (szMode = SZ_BEST_SPEED only): such a file is
 * refused -- a message, *oriData = NULL and 0 bytes -- instead of being decoded. */
unsigned long sz_lossless_decompress65536bytes(int losslessCompressor, unsigned char *compressBytes, unsigned long cmpSize, unsigned char **oriData) {
    (void)losslessCompressor; (void)compressBytes; (void)cmpSize;
    printf("Error: this file was written with a zstd/gzip pass (szMode != SZ_BEST_SPEED); libSZ_b200 reads SZ_BEST_SPEED streams only.\n");
    *oriData = NULL;                     /* nothing decoded: 0 bytes, no buffer -- the caller decides (a library does not end the process) */
    return 0;
}
