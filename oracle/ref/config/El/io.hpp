/* stand-in for the reference's <El/io.hpp>, which src/blas_like/level3/Trrk/Local.hpp includes for DimsString only
 * (oracle/ref/ref_summa.cpp supplies that); the real header needs the whole library */
