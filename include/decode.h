/*
 * decode.h -- drop-in for the reference's src/decode.h:35-41 (same symbol, same arguments).
 *
 *   inFile    original file name: "<inFile>.METADATA" is read; the output overwrites inFile when
 *             outFile is NULL (src/decode.cu:410-417)
 *   confFile  k whitespace-separated fragment file names; line i supplies row i of the decode input
 *             and its index is atoi(name + 1) (src/decode.cu:293-319)
 *   outFile   where the first totalSize reconstructed bytes go (may be NULL)
 *   gridDimXSize, streamNum   as in encode.h
 *
 * METADATA: both the GPU writer's form (size, "p k", then the (k+p) x k matrix, src/encode.cu:61-101)
 * and the CPU program's two-line form (src/cpu-rs.c:465-476; matrix regenerated) are accepted.
 * The k x k inversion runs on the GPU (single-block Gauss-Jordan with row pivoting); a survivor set
 * that cannot reconstruct the data prints "Cannot decode: ..." and exits with status 1 instead of
 * silently writing wrong data as the reference does.
 */
#ifndef RS_DECODE_H_
#define RS_DECODE_H_
#ifdef __cplusplus
extern "C" {
#endif
void decode_file(char *inFile, char *confFile, char *outFile, int gridDimXSize, int streamNum);
#ifdef __cplusplus
}
#endif
#endif
