/* One binary, both curves: links libb200prover.so and libb200prover_bls12377.so together (see smoke.c). */
#include <stdio.h>
#include <string.h>
int smoke_bn254(const char *path);
int smoke_bls12377(const char *path);
int main(int argc, char **argv) {
    if (argc != 3) { fprintf(stderr, "usage: %s bn254|bls12377 input-file\n", argv[0]); return 1; }
    return strcmp(argv[1], "bls12377") == 0 ? smoke_bls12377(argv[2]) : smoke_bn254(argv[2]);
}
