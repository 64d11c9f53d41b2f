// Streaming VBD colour pass instantiated for cloth with wedge records (streams.cuh: ST_WEDGE -- a fan triangle's StVK
// membrane and the two discrete-bending hinges that share its normal as one record) next to the three single-incidence
// streams that carry what completes no wedge; anchors through the CSR path.
#include "vbd_stream.cuh"

namespace spgx
{
void launchStreamClothWedge(const VbdArgs &args, int lanes, int occ, int begin, int end, int sliceBase, double invdtSquared,
                            cudaStream_t stream)
{
    launchStream<kStreamsCloth, kKindsAnchor>(args, lanes, occ, begin, end, sliceBase, invdtSquared, stream);
}
}  // namespace spgx
