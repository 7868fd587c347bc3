/*
 * lda - the reference's command line on top of libonilda (B200 kernels).
 *
 *   lda est [alpha] [k] [settings] [#processes] [data] [random/seeded/<model root>] [output directory]
 *   lda inf [settings] [model root] [data] [output name]
 *
 * Same argv positions, same files in and out as oni-ml/oni-lda-c/lda-estimate.c:403-468,
 * so OniLDACWrapper.scala:83-84 can exec it unchanged (also under `mpiexec -n P`:
 * when an MPI launcher starts several copies, only the copy with rank 0 in
 * PMI_RANK / OMPI_COMM_WORLD_RANK works, the others exit 0 at once).
 *
 * [#processes] is the number of GPUs of this box to shard the documents over
 * (clamped to the devices present; results do not depend on it beyond
 * floating-point summation order).  The RNG seed of random/seeded comes from
 * time() as in the reference (:414-416) unless ONILDA_SEED is set.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <time.h>

#include "onilda.h"

static int launcher_rank(void)
{
    const char* names[] = { "PMI_RANK", "OMPI_COMM_WORLD_RANK", "PMIX_RANK", "MV2_COMM_WORLD_RANK" };
    for (unsigned i = 0; i < sizeof names / sizeof names[0]; i++) {
        const char* v = getenv(names[i]);
        if (v) return atoi(v);
    }
    return 0;
}

static int die(const char* what)
{
    fprintf(stderr, "lda: %s: %s\n", what, onilda_last_error());
    return 1;
}

/* CUDA context creation runs beside the parse of model.dat (both take about half a second at cfg 2) */
static void* warmup_thread(void* arg)
{
    onilda_device_warmup(*(int*)arg);
    return NULL;
}

int main(int argc, char* argv[])
{
    if (argc <= 1) {
        printf("usage : lda est [alpha] [k] [settings] [#processes] [data] [random/seeded/*] [output directory]\n");
        printf("        lda inf [settings] [model] [data] [output filename]\n");
        return 0;
    }
    if (launcher_rank() != 0) return 0;

    uint32_t seed = (uint32_t)time(NULL);
    const char* s = getenv("ONILDA_SEED");
    if (s) seed = (uint32_t)strtoul(s, NULL, 10);

    onilda_settings settings;
    onilda_corpus* corpus = NULL;
    if (strcmp(argv[1], "est") == 0) {
        if (argc < 9) { fprintf(stderr, "lda: est needs 7 arguments\n"); return 2; }
        float alpha = (float)atof(argv[2]);      /* float, as lda-estimate.c:425 */
        int ntopics = atoi(argv[3]);
        int ndev = atoi(argv[5]);
        if (onilda_settings_read(argv[4], &settings) != 0) return die("settings");
        struct timespec t0, t1;
        pthread_t warm;
        int warm_ndev = ndev;
        const int warming = pthread_create(&warm, NULL, warmup_thread, &warm_ndev) == 0;
        clock_gettime(CLOCK_MONOTONIC, &t0);
        const int parse_rc = onilda_corpus_from_dat(argv[6], &corpus);
        clock_gettime(CLOCK_MONOTONIC, &t1);
        if (warming) pthread_join(warm, NULL);     /* a failure there is reported by run_em with its own message */
        if (parse_rc != 0) return die("corpus");
        if (getenv("ONILDA_TIMING"))
            fprintf(stderr, "[timing] read %s: %.3f s\n", argv[6], (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec));
        int64_t D, N; int32_t V;
        onilda_corpus_dims(corpus, &D, &N, &V, NULL, NULL);
        printf("number of docs    : %lld\n", (long long)D);
        printf("number of terms   : %d\n", V);
        int rc = onilda_run_em(argv[7], argv[8], corpus, ntopics, (double)alpha, &settings, ndev, seed);
        onilda_corpus_free(corpus);
        if (rc != 0) return die("est");
        return 0;
    }
    if (strcmp(argv[1], "inf") == 0) {
        if (argc < 6) { fprintf(stderr, "lda: inf needs 4 arguments\n"); return 2; }
        if (onilda_settings_read(argv[2], &settings) != 0) return die("settings");
        if (onilda_corpus_from_dat(argv[4], &corpus) != 0) return die("corpus");
        int rc = onilda_infer_files(argv[3], argv[5], corpus, &settings, -1);
        onilda_corpus_free(corpus);
        if (rc != 0) return die("inf");
        return 0;
    }
    return 0;
}
