/* oracle/ref_main.cpp — TEST INFRASTRUCTURE. Two-line entry point linked with the
 * UNMODIFIED reference sources (src/gmove.cpp etc.) so that `poregen_ref gmove …`
 * behaves like `poregen gmove …` (reference src/main.c:76 calls gmove(argc-1, argv+1)).
 * The reference's own main.c is not used because it links every sub-tool
 * (reform/realign need real htslib). */
#include <stdio.h>
#include <string.h>
int gmove(int argc, char *argv[]);
int main(int argc, char *argv[]) {
    if (argc < 2 || strcmp(argv[1], "gmove") != 0) {
        fprintf(stderr, "usage: poregen_ref gmove [options] reads.blow5 alignment output_dir\n");
        return 1;
    }
    return gmove(argc - 1, argv + 1);
}
