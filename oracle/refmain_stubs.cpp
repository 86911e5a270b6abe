/* oracle/refmain_stubs.cpp - TEST INFRASTRUCTURE. What the reference's UNMODIFIED src/main.c needs at link time besides
 * gmove() and reform() (which come from this build's libporegen_gmove_cli.so, include/pgm_gmove_cli.h): the log-level
 * accessors its ERROR macro expands to (reference src/error.h:16,29-30; defined in src/poregen.cpp:240-246, a file that
 * drags in the whole batch framework) and the three sub-tools that are not on the gmove path. */
#include <stdio.h>
#include "error.h"   /* the reference's own header, from -I $REF/src */

static enum poregen_log_level_opt g_level = LOG_VERB;
enum poregen_log_level_opt get_log_level() { return g_level; }
void set_log_level(enum poregen_log_level_opt level) { g_level = level; }

static int not_built(const char *what) {
    fprintf(stderr, "[poregen] %s is not part of the gmove hot-path build\n", what);
    return 1;
}
int subtool0(int, char *[]) { return not_built("subtool0"); }
int kmer_freq(int, char *[]) { return not_built("kmer_freq"); }
int realign(int, char *[]) { return not_built("realign"); }
