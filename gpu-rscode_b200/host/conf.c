/* conf.c -- the decode configuration file: k whitespace-separated fragment file names; the name on
 * position i feeds row i of the decode input and its fragment index is atoi(name + 1), i.e. the
 * digits after the leading '_' (reference src/decode.cu:302-305, src/cpu-rs.c:635-638). */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "rs_host.h"

int rs_read_conf(const char *path, int k, rs_conf *conf)
{
    conf->k = 0;
    conf->name = NULL;
    conf->index = NULL;
    FILE *fp = fopen(path, "r");
    if (!fp)
        return -1;
    conf->name = calloc((size_t)k, sizeof(char *));
    conf->index = calloc((size_t)k, sizeof(int));
    char buf[4096];
    int rc = 0;
    for (int i = 0; i < k; ++i) {
        if (fscanf(fp, "%4095s", buf) != 1) {
            rc = -2;
            break;
        }
        conf->name[i] = strdup(buf);
        conf->index[i] = atoi(buf + 1);
        conf->k = i + 1;
    }
    fclose(fp);
    if (rc)
        rs_free_conf(conf);
    return rc;
}

void rs_free_conf(rs_conf *conf)
{
    if (conf->name)
        for (int i = 0; i < conf->k; ++i)
            free(conf->name[i]);
    free(conf->name);
    free(conf->index);
    conf->name = NULL;
    conf->index = NULL;
    conf->k = 0;
}
