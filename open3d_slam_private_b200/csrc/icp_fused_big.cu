// The persistent registration kernel for k-best lists of 32 and 64 entries (17 <= knn <= 64): the same source as
// icp_fused.cu, compiled as a second translation unit so that the five instantiations build in two parallel halves.
#define B200_ICP_FUSED_PART 1
#include "icp_fused.cu"
