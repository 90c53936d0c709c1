/*
 * sz.h -- drop-in public C API of the B200-native SZ2 hot path (libSZ_b200.so).
 *
 * Same names, argument meaning, return codes, struct layouts and exported globals as the reference's
 * sz/include/sz.h (cited per item as "ref:"), restricted to the path BASELINE.json's north_star names:
 * SZ_FLOAT / SZ_DOUBLE, 2D/3D/4D, errorBoundMode ABS / REL / ABS_AND_REL / ABS_OR_REL, szMode = SZ_BEST_SPEED,
 * withLinearRegression = YES.  Everything else returns an error (SZ_NSCS / NULL) -- there is NO CPU fallback:
 * every call that produces or consumes a stream runs the CUDA kernels on the current device.
 *
 * Pointers are HOST pointers, exactly as in the reference; see include/szb200.h for the device-pointer
 * entry points used by the benchmark and by multi-GPU slab sharding.
 */
#ifndef SZ_B200_SZ_H
#define SZ_B200_SZ_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ref: sz/include/defines.h:14-16 */
#define SZ_VER_MAJOR 2
#define SZ_VER_MINOR 1
#define SZ_VER_BUILD 12
#define SZ_VER_REVISION 0

/* ref: sz/include/defines.h:26-41 (error-bound modes) */
#define ABS 0
#define REL 1
#define VR_REL 1
#define ABS_AND_REL 2
#define ABS_OR_REL 3
#define PSNR 4
#define NORM 5
#define PW_REL 10
#define ABS_AND_PW_REL 11
#define ABS_OR_PW_REL 12
#define REL_AND_PW_REL 13
#define REL_OR_PW_REL 14

/* ref: sz/include/defines.h:43-52 */
#define SZ_FLOAT 0
#define SZ_DOUBLE 1
#define SZ_UINT8 2
#define SZ_INT8 3
#define SZ_UINT16 4
#define SZ_INT16 5
#define SZ_UINT32 6
#define SZ_INT32 7
#define SZ_UINT64 8
#define SZ_INT64 9

#define LITTLE_ENDIAN_DATA 0
#define BIG_ENDIAN_DATA 1
#define LITTLE_ENDIAN_SYSTEM 0
#define BIG_ENDIAN_SYSTEM 1

/* ref: sz/include/defines.h:65-71 */
#define SZ_BEST_SPEED 0
#define SZ_BEST_COMPRESSION 1
#define SZ_DEFAULT_COMPRESSION 2
#define SZ_TEMPORAL_COMPRESSION 3
#define SZ_NO_REGRESSION 0
#define SZ_WITH_LINEAR_REGRESSION 1

/* ref: sz/include/defines.h:84-90 */
#define SZ_SCES 0
#define SZ_NSCS -1
#define SZ_FERR -2
#define SZ_TERR -3
#define SZ_DERR -4
#define SZ_MERR -5
#define SZ_BERR -6

#define MetaDataByteLength 28         /* ref: defines.h:97 */
#define MetaDataByteLength_double 36  /* ref: defines.h:98 */
#define MIN_NUM_OF_ELEMENTS 20        /* ref: sz/include/sz.h:81 */

#define SZ 101                        /* sol_ID, ref: sz/include/defines.h:19 */
#define SZ_Transpose 104
#define GZIP_COMPRESSOR 0
#define ZSTD_COMPRESSOR 1

/* ref: sz/include/sz.h:164-198 -- field order and types are part of the ABI (callers memcpy this struct) */
typedef struct sz_params
{
	int dataType;
	unsigned int max_quant_intervals;
	unsigned int quantization_intervals;
	unsigned int maxRangeRadius;
	int sol_ID;
	int losslessCompressor;
	int sampleDistance;
	float predThreshold;
	int szMode;
	int gzipMode;
	int errorBoundMode;
	double absErrBound;
	double relBoundRatio;
	double psnr;
	double normErr;
	double pw_relBoundRatio;
	int segment_size;
	int pwr_type;

	int protectValueRange;
	float fmin, fmax;
	double dmin, dmax;

	int snapshotCmprStep;
	int predictionMode;

	int accelerate_pw_rel_compression;
	int plus_bits;

	int randomAccess;
	int withRegression;
} sz_params;

/* ref: sz/include/sz.h:200-209 */
typedef struct sz_metadata
{
	int versionNumber[3];
	int isConstant;
	int isLossless;
	int sizeType;
	size_t dataSeriesLength;
	int defactoNBBins;
	struct sz_params* conf_params;
} sz_metadata;

/* ref: sz/include/sz.h:211-217 */
typedef struct sz_exedata
{
	char optQuantMode;
	int intvCapacity;
	int intvRadius;
	unsigned int SZ_SIZE_TYPE;
} sz_exedata;

/* ref: sz/include/sz.h:232-240 -- exported data globals, read and written directly by callers (example/sz.c:317-338) */
extern int versionNumber[4];
extern int dataEndianType;
extern int sysEndianType;
extern sz_params *confparams_cpr;
extern sz_params *confparams_dec;
extern sz_exedata *exe_params;

/* ref: sz/src/sz.c:60 (SZ_Init), :75 (SZ_Init_Params), sz/src/conf.c:74-461 (config keys and defaults) */
int SZ_Init(const char *configFilePath);
int SZ_Init_Params(sz_params *params);
/* ref: sz/src/sz.c:925 */
void SZ_Finalize(void);

/* ref: sz/src/sz.c:103-160, :162-282 */
size_t computeDataLength(size_t r5, size_t r4, size_t r3, size_t r2, size_t r1);
int computeDimension(size_t r5, size_t r4, size_t r3, size_t r2, size_t r1);
int filterDimension(size_t r5, size_t r4, size_t r3, size_t r2, size_t r1, size_t* correctedDimension);

/* ref: sz/src/sz.c:436 -- uses confparams_cpr->{errorBoundMode,absErrBound,relBoundRatio} */
unsigned char *SZ_compress(int dataType, void *data, size_t *outSize, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1);
/* ref: sz/src/sz.c:294-391 -- returns a malloc'd stream (caller frees) or NULL; r1 is the fastest dimension, unused dims are 0 */
unsigned char* SZ_compress_args(int dataType, void *data, size_t *outSize, int errBoundMode, double absErrBound,
double relBoundRatio, double pwrBoundRatio, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1);
/* ref: sz/src/sz.c:393-408 -- writes into caller memory */
int SZ_compress_args2(int dataType, void *data, unsigned char* compressed_bytes, size_t *outSize,
int errBoundMode, double absErrBound, double relBoundRatio, double pwrBoundRatio,
size_t r5, size_t r4, size_t r3, size_t r2, size_t r1);

/* ref: sz/src/sz.c:486-577 -- returns a malloc'd array of N elements (caller frees) or NULL */
void *SZ_decompress(int dataType, unsigned char *bytes, size_t byteLength, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1);
/* ref: sz/src/sz.c:584-680 -- copies into caller memory, returns the element count */
size_t SZ_decompress_args(int dataType, unsigned char *bytes, size_t byteLength, void* decompressed_array, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1);

/* ref: sz/src/sz.c:683-923 */
sz_metadata* SZ_getMetadata(unsigned char* bytes);
void SZ_printMetadata(sz_metadata* metadata);

/* ref: sz/src/ByteToolkit.c:874-972, :974-1063 (stream params block) */
void convertSZParamsToBytes(sz_params* params, unsigned char* result);
void convertBytesToSZParams(unsigned char* bytes, sz_params* params);

/* ref: sz/src/sz.c:1362-1510 -- compressor names "SZ", "SZ2.0", "SZ2.1"; userPara is a sz_params* (or NULL) */
unsigned char* SZ_compress_customize(const char* appName, void* userPara, int dataType, void* data, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1, size_t *outSize, int *status);
unsigned char* SZ_compress_customize_threadsafe(const char* cmprName, void* userPara, int dataType, void* data, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1, size_t *outSize, int *status);
void* SZ_decompress_customize(const char* appName, void* userPara, int dataType, unsigned char* bytes, size_t byteLength, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1, int* status);
void* SZ_decompress_customize_threadsafe(const char* cmprName, void* userPara, int dataType, unsigned char* bytes, size_t byteLength, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1, int *status);

/* ---- file helpers used by the reference's example programs (ref: sz/include/rw.h:43-70), implemented in sz_io.c ---- */
unsigned char *readByteData(char *srcFilePath, size_t *byteLength, int *status);
double *readDoubleData(char *srcFilePath, size_t *nbEle, int *status);
float *readFloatData(char *srcFilePath, size_t *nbEle, int *status);
void writeByteData(unsigned char *bytes, size_t byteLength, char *tgtFilePath, int *status);
void writeFloatData_inBytes(float *data, size_t nbEle, char *tgtFilePath, int *status);
void writeDoubleData_inBytes(double *data, size_t nbEle, char *tgtFilePath, int *status);
/* ---- what the reference's command-line tool (example/sz.c) needs beyond those (ref: sz/include/rw.h:31,66,67,78 and
 *      sz/include/utility.h:34-38), also in sz_io.c.  A file that carries a zstd/gzip pass (szMode != SZ_BEST_SPEED) is recognised
 *      and refused: that back-end is outside this library. ---- */
int checkFileExistance(char *filePath);
void writeStrings(int nbStr, char *str[], char *tgtFilePath, int *status);
void writeFloatData(float *data, size_t nbEle, char *tgtFilePath, int *status);
void writeDoubleData(double *data, size_t nbEle, char *tgtFilePath, int *status);
void *detransposeData(void *data, int dataType, size_t r5, size_t r4, size_t r3, size_t r2, size_t r1);
int is_lossless_compressed_data(unsigned char *compressedBytes, size_t cmpSize);
unsigned long sz_lossless_decompress65536bytes(int losslessCompressor, unsigned char *compressBytes, unsigned long cmpSize, unsigned char **oriData);

#ifdef __cplusplus
}
#endif
#endif
