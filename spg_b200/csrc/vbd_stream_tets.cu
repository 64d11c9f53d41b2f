// Streaming VBD colour pass instantiated for tetrahedral meshes: stable Neo-Hookean / StVK tet streams, anchors through
// the CSR path.
#include "vbd_stream.cuh"

namespace spgx
{
void launchStreamTets(const VbdArgs &args, int lanes, int occ, int begin, int end, int sliceBase, double invdtSquared,
                      cudaStream_t stream)
{
    launchStream<kStreamsTets, kKindsAnchor>(args, lanes, occ, begin, end, sliceBase, invdtSquared, stream);
}
}  // namespace spgx
