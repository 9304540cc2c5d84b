/* Stand-in so that the reference's zklaim/zklaim.h (which includes zklaim_ecc.h ->
 * <gcrypt.h>) parses in the oracle build.  No libgcrypt function is called on the
 * Groth16 prove path.  TEST INFRASTRUCTURE only. */
#ifndef ZKB_ORACLE_GCRYPT_SHIM_H
#define ZKB_ORACLE_GCRYPT_SHIM_H
#include <stddef.h>
typedef struct gcry_sexp *gcry_sexp_t;
typedef struct gcry_mpi *gcry_mpi_t;
typedef unsigned int gcry_error_t;
#endif
