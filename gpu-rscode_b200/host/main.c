/* main.c -- the RS command line, same flags and rules as the reference's src/main.c:47-164:
 *   encode:  RS -k K -n N -e file            [-p maxGridDimX] [-s streams]
 *   decode:  RS -d -i file -c conf [-o out]  [-p maxGridDimX] [-s streams]
 * -d must come before -i/-c/-o (otherwise the usage text is shown, src/main.c:104-137); decoding
 * takes k and n from the METADATA file.  Two deliberate differences: without -s the stream count is
 * chosen by the engine (the reference defaults to 1, which never overlaps copies and kernels); and the
 * documented upper-case spellings -K -N -E -I -C -O -S -P take their argument here -- the reference's
 * getopt string "Ss:Pp:Kk:..." gives them none, so it dereferences a NULL optarg and crashes.
 */
#include <assert.h>
#include <getopt.h>
#include <stdio.h>
#include <stdlib.h>

#include "decode.h"
#include "encode.h"
#include "rs_host.h"

static void show_help_info(void)
{
    printf("Usage:\n");
    printf("[-h]: show usage information\n");
    printf("Encode: [-k|-K nativeBlockNum] [-n|-N totalBlockNum] [-e|-E fileName]\n");
    printf("Decode: [-d|-D] [-k|-K nativeBlockNum] [-n|-N totalBlockNum] \n\t [-i|-I originalFileName] [-c|-C config] [-o|-O output]\n");
    printf("For encoding, the -k, -n, and -e options are all necessary.\n");
    printf("For decoding, the -d, -i, and -c options are all necessary.\n");
    printf("If the -o option is not set, the original file name will be chosen as the output file name by default.\n");
    printf("Performance-tuning Options:\n");
    printf("[-p|-P]: set maxmimum blockDimX\n");
    printf("[-s|-S]: set stream number\n");
    exit(0);
}

int main(int argc, char *argv[])
{
    int k = 0, n = 0, grid = 0, streams = 0;
    char *inFile = NULL, *confFile = NULL, *outFile = NULL;
    enum { NONE, ENCODE, DECODE } op = NONE;
    int option;
    while ((option = getopt(argc, argv, "S:s:P:p:K:k:N:n:E:e:I:i:C:c:O:o:Ddh")) != -1) {
        switch (option) {
        case 'S': case 's': streams = atoi(optarg); break;
        case 'P': case 'p': grid = atoi(optarg); break;
        case 'K': case 'k': k = atoi(optarg); break;
        case 'N': case 'n': n = atoi(optarg); break;
        case 'E': case 'e': inFile = optarg; op = ENCODE; break;
        case 'D': case 'd': op = DECODE; break;
        case 'I': case 'i':
            if (op != DECODE) show_help_info();
            inFile = optarg;
            break;
        case 'C': case 'c':
            if (op != DECODE) show_help_info();
            confFile = optarg;
            break;
        case 'O': case 'o':
            if (op != DECODE) show_help_info();
            outFile = optarg;
            break;
        default: show_help_info(); break;
        }
    }
    switch (op) {
    case ENCODE:
        assert(k != 0);
        assert(n != 0);
        rs_cli_prepare_gpus(inFile, 0); /* our own process: start CUDA on the GPUs this job will use, not on all */
        encode_file(inFile, k, n - k, grid, streams);
        break;
    case DECODE:
        assert(inFile != NULL);
        assert(confFile != NULL);
        rs_cli_prepare_gpus(inFile, 1);
        decode_file(inFile, confFile, outFile, grid, streams);
        break;
    default:
        break;
    }
    return 0;
}
