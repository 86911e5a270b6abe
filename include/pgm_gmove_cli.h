/* pgm_gmove_cli.h - the link-level boundary of the B200 build: the two sub-tool entry points the reference's
 * dispatcher calls (reference src/main.c:43,47 declarations; :76,80 calls), exported by libporegen_gmove_cli.so.
 *
 * The reference compiles main.c as C++ (Makefile:5,39 `-x c++`), so its main.o wants the C++-mangled symbols
 * _Z5gmoveiPPc and _Z6reformiPPc; both are exported under those names, and under unmangled C names for any other
 * caller. argv[0] is the sub-tool's own name ("gmove" / "reform"), exactly what main() passes (argc-1, argv+1).
 * Exit codes, messages, option table and output files are the reference's (src/gmove.cpp:49-72, 240-537;
 * src/reform.cpp:79-371). tests/test_cli_link.py links the reference's UNMODIFIED main.c against this library.
 */
#ifndef PGM_GMOVE_CLI_H
#define PGM_GMOVE_CLI_H

#ifdef __cplusplus
int gmove(int argc, char *argv[]);   /* replaces reference src/gmove.cpp:213 */
int reform(int argc, char *argv[]);  /* replaces reference src/reform.cpp:79 */
extern "C" {
#endif
int pgm_gmove_main(int argc, char **argv);   /* = gmove(argc, argv) */
int pgm_reform_main(int argc, char **argv);  /* = reform(argc, argv) */
/* apply_transformation <input_file> <output_file> <A> <B> <C> <D> of the workflow script (reference
 * scripts/poregen.sh:87-129): argv[0] is ignored, argv[1..6] are the six arguments as TEXT (bc's result scales depend on
 * the digits written). Host text work, no GPU. */
int pgm_transform_main(int argc, char **argv);
const char *pgm_cli_build_info(void);        /* sha256 of the command line's sources at build time */
#ifdef __cplusplus
}
#endif
#endif /* PGM_GMOVE_CLI_H */
