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
