/* empty stand-in: src/blas_like/level3/Gemm/NN.hpp of this fork includes this header at file scope; nothing of it is needed by the SUMMA loops compiled in oracle/ref/ref_summa.cpp */
