// Compile check of adapter/b200_funcs_shim.h against an UNMODIFIED reference client that is a
// header: gpu/elementwise_grad.h.  Instantiates its reduce path (ReduceWrapper -> phi::SumKernel),
// its one- and two-output broadcast paths (GetGradXOrYOut / GetGradXAndYOut with the Max/MinGrad
// functors, as ElementwiseMaxGrad does) and the InverseFunctor reduce of default_elementwise_sub_grad (:258-308).
#include "paddle/phi/kernels/gpu/elementwise_grad.h"

#include "paddle/phi/kernels/funcs/elementwise_functor.h"

namespace phi {
template void ReduceWrapper<float>(const GPUContext&, int, DenseTensor*, DenseTensor*);
template void GetGradXOrYOut<float, funcs::MaxGradXFunctor<float>>(const GPUContext&, const Place&, int,
                                                                    std::vector<const DenseTensor*>,
                                                                    const DenseTensor&, DenseTensor*,
                                                                    funcs::MaxGradXFunctor<float>);
template void GetGradXAndYOut<float, funcs::MaxGradXYFunctor<float, float>>(
    const GPUContext&, const Place&, int, std::vector<const DenseTensor*>, const DenseTensor&, DenseTensor*,
    DenseTensor*, funcs::MaxGradXYFunctor<float, float>);
template void DefaultElementwiseAddGrad<float>(const GPUContext&, const DenseTensor&, const DenseTensor&,
                                               const DenseTensor&, const DenseTensor&, DenseTensor*,
                                               DenseTensor*, int);
template void default_elementwise_sub_grad<float>(const GPUContext&, const DenseTensor&, const DenseTensor&,
                                                  const DenseTensor&, const DenseTensor&, DenseTensor*,
                                                  DenseTensor*, int);
template void default_elementwise_sub_grad<phi::dtype::bfloat16>(const GPUContext&, const DenseTensor&,
                                                                 const DenseTensor&, const DenseTensor&,
                                                                 const DenseTensor&, DenseTensor*,
                                                                 DenseTensor*, int);
}  // namespace phi
