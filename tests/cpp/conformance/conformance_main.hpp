// main() of the conformance binaries: brings the GPU up (unless the suite is host
// only), runs every registered test or those whose name contains argv[1].
#ifndef RAILS_CONFORMANCE_MAIN_HPP
#define RAILS_CONFORMANCE_MAIN_HPP

#ifndef RAILS_CONFORMANCE_NO_GPU
#include "rails_cuda.h"
#endif

#include <cstdlib>

int main(int argc, char **argv)
{
#ifndef RAILS_CONFORMANCE_NO_GPU
    if (rails_cuda_init(0))
    {
        std::cerr << "rails_cuda_init failed: " << rails_cuda_last_error() << std::endl;
        return 2;
    }
#endif
    std::srand(1);
    int rc = RAILS_RUN_ALL_TESTS(argc > 1 ? argv[1] : nullptr);
#ifndef RAILS_CONFORMANCE_NO_GPU
    rails_cuda_finalize();
#endif
    return rc;
}

#endif
