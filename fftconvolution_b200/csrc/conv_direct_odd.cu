// Odd-width instantiations of the bulk-staged direct kernel (conv_direct_bulk.cuh), in their own translation unit so
// that they compile beside conv_direct.cu.
#include "conv_direct_bulk.cuh"

namespace b2c {

direct_fn direct_bulk_variant_odd(int hk, int nbuf, int ti) { return direct_bulk_variant_o<true>(hk, nbuf, ti); }

}  // namespace b2c
