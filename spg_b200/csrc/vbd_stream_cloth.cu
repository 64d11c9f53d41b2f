// Streaming VBD colour pass instantiated for cloth: discrete bending (two canonical streams) + StVK membrane streams,
// anchors through the CSR path.  Groups whose incidences were joined into wedge records run vbd_stream_cloth_wedge.cu.
#include "vbd_stream.cuh"

namespace spgx
{
void launchStreamCloth(const VbdArgs &args, int lanes, int occ, int begin, int end, int sliceBase, double invdtSquared,
                       cudaStream_t stream)
{
    launchStream<kStreamsClothSingle, kKindsAnchor>(args, lanes, occ, begin, end, sliceBase, invdtSquared, stream);
}
}  // namespace spgx
