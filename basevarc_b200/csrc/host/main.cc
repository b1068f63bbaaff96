// main.cc -- the `BaseVarC`-compatible command: only `basetype` (with complete batch files) is served here;
// `--load`, popmatrix and concat stay with the reference binary (/root/reference/src/BaseVarC.cpp:151-174).
#include <stdio.h>
#include <string.h>

#include "../../../include/bvc_host.h"

int main(int argc, char** argv)
{
    if (argc > 1 && strcmp(argv[1], "basetype") == 0) return bvc_basetype_main(argc - 1, argv + 1);
    fprintf(stderr, "Usage  : BaseVarC_b200 basetype --rerun [options]   (GPU calling core of BaseVarC basetype)\n"
                    "         popmatrix / concat / basetype --load: use the reference BaseVarC binary\n");
    return 0;
}
