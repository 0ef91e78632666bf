/*
 * encode.h -- drop-in for the reference's src/encode.h:33-39 (same symbol, same arguments).
 *
 *   fileName        file to encode; fragments "_<i>_<fileName>" and "<fileName>.METADATA" are written
 *                   to the current directory (src/encode.cu:283-285,434-465)
 *   nativeBlockNum  k, parityBlockNum  p = n - k
 *   gridDimXSize    the reference's -p: cap on gridDim.x of the multiply kernel (<= 0 or too large:
 *                   prints "Valid grid size: (0, %d]" and uses the default, src/encode.cu:348-355)
 *   streamNum       the reference's -s: number of pipeline slots (streams) per GPU
 *
 * Behaviour kept: blocking call, IO errors print the reference's message and exit(0), the "Total GPU
 * encoding time using multiple devices" line, no exchange between GPUs.  How the work is spread differs from
 * the reference (one pthread per visible GPU, each a contiguous byte range, src/encode.cu:357-403): the
 * column space is cut into slabs that go to whichever GPU worker is free, and a job starts one GPU worker
 * per 32 GiB of input (RS_GPUS=<n> overrides) because every further CUDA context costs start-up time.  Changed on purpose: 64-bit sizes (files >= 2 GiB work), the
 * tail of the last native fragment is zero-padded as in cpu-rs.c:502 (the GPU reference leaves
 * malloc garbage there, src/encode.cu:325), no cudaDeviceReset on return.
 */
#ifndef RS_ENCODE_H_
#define RS_ENCODE_H_
#ifdef __cplusplus
extern "C" {
#endif
void encode_file(char *fileName, int nativeBlockNum, int parityBlockNum, int gridDimXSize, int streamNum);
#ifdef __cplusplus
}
#endif
#endif
