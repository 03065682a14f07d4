// Host-side unit test of the functor classification in adapter/b200_funcs_shim.h: the kps:: unary
// functors, functors shaped like the client-local ones of gpu/p_norm_kernel.cu:25-52, and a
// client-local reducer like BitwiseAdd (kps/compare_kernel.cu:36-45).  Built and run by
// tests/test_adapter_compile.py; prints "probe ok".
#include <cstdio>

#include "paddle/phi/kernels/funcs/reduce_function.h"

namespace client {
template <typename T>
struct NonzeroFunctor {
  HOSTDEVICE inline T operator()(const T x) const { return static_cast<T>(static_cast<double>(x) != 0); }
};
template <typename T>
struct AbsFunctor {
  HOSTDEVICE inline T operator()(const T x) const { return x < T(0) ? -x : x; }
};
template <typename T>
struct UnsignedPowFunctor {
  HOSTDEVICE explicit inline UnsignedPowFunctor(float p) : porder(p) {}
  HOSTDEVICE inline T operator()(const T x) const {
    return static_cast<T>(pow(static_cast<double>(x < T(0) ? -x : x), static_cast<double>(porder)));
  }
  float porder;
};
template <typename T>
struct CubeFunctor {
  HOSTDEVICE inline T operator()(const T x) const { return x * x * x; }
};
template <typename T>
struct BitwiseAdd {
  inline T initial() { return static_cast<T>(true); }
  __host__ __device__ __forceinline__ T operator()(const T& a, const T& b) const { return a & b; }
};
}  // namespace client

#define EXPECT(cond)                                       \
  if (!(cond)) {                                           \
    printf("FAILED %s:%d %s\n", __FILE__, __LINE__, #cond); \
    return 1;                                              \
  }

int main() {
  using namespace phi::b200shim;  // NOLINT
  namespace kps = phi::kps;
  EXPECT(ClassifyUnary<float>(kps::IdentityFunctor<float>()).cls == kIdentity);
  EXPECT((ClassifyUnary<phi::dtype::bfloat16>(kps::IdentityFunctor<phi::dtype::bfloat16, float>()).cls == kIdentity));
  EXPECT(ClassifyUnary<int64_t>(kps::IdentityFunctor<int64_t>()).cls == kIdentity);
  EXPECT(ClassifyUnary<bool>(kps::IdentityFunctor<bool>()).cls == kIdentity);
  EXPECT(ClassifyUnary<double>(kps::InverseFunctor<double>()).cls == kNegate);
  EXPECT(ClassifyUnary<float>(kps::SquareFunctor<float>()).cls == kSquare);
  const UnaryInfo d = ClassifyUnary<float>(kps::DivideFunctor<float>(7));
  EXPECT(d.cls == kScale && d.param > 0.1428 && d.param < 0.1429);
  const UnaryInfo dh = ClassifyUnary<phi::dtype::float16>(kps::DivideFunctor<phi::dtype::float16>(1000));
  EXPECT(dh.cls == kScale && dh.param > 0.00098 && dh.param < 0.00102);
  EXPECT(ClassifyUnary<float>(client::NonzeroFunctor<float>()).cls == kNonzero);
  EXPECT(ClassifyUnary<double>(client::AbsFunctor<double>()).cls == kAbs);
  const UnaryInfo p = ClassifyUnary<float>(client::UnsignedPowFunctor<float>(2.5f));
  EXPECT(p.cls == kPowAbs && p.param > 2.499 && p.param < 2.501);
  const UnaryInfo q = ClassifyUnary<float>(client::UnsignedPowFunctor<float>(1.0f / 3.0f));
  EXPECT(q.cls == kPowAbs && q.param > 0.3332 && q.param < 0.3335);
  EXPECT(ClassifyUnary<float>(client::CubeFunctor<float>()).cls == kUnknownUnary);
  EXPECT((ReduceOpId<kps::AddFunctor, float>() == B200PHI_SUM));
  EXPECT((ReduceOpId<kps::MaxFunctor, double>() == B200PHI_MAX));
  EXPECT((ReduceOpId<kps::MinFunctor, int64_t>() == B200PHI_MIN));
  EXPECT((ReduceOpId<kps::MulFunctor, float>() == B200PHI_PROD));
  EXPECT((ReduceOpId<kps::LogicalOrFunctor, bool>() == B200PHI_ANY));
  EXPECT((ReduceOpId<client::BitwiseAdd, bool>() == B200PHI_ALL));
  printf("probe ok\n");
  return 0;
}
