// Functor plug-ins for the broadcast engine (b200phi_map): compare / logical / bitwise,
// the ternary functors of the backward kernels, and the unary activation set.
// Each one restates a reference functor; the op id is a template argument so a kernel
// holds exactly one of them.
//   funcs/compare_functors.h:20-60     funcs/logical_functor.h:20-42
//   funcs/bitwise_functors.h:20-48     funcs/elementwise_functor.h:173-179,478-531
//   funcs/activation_functor.h (Cuda*Functor forms)   gpu/p_norm_kernel.cu:26-52
#pragma once

#include <limits>

#include "../../include/b200phi.h"
#include "prim.cuh"

namespace b200phi {

constexpr bool map_out_is_bool(int op) { return op < B200PHI_MAP_BITWISE_AND; }  // compare, logical
constexpr int map_arity(int op) {
  if (op == B200PHI_MAP_LOGICAL_NOT || op == B200PHI_MAP_BITWISE_NOT) return 1;
  if (op >= B200PHI_MAP_EQUAL_MUL && op <= B200PHI_MAP_DIV_GRAD_Y) return 3;
  if (op >= B200PHI_MAP_EXP) return 1;
  return 2;
}

// EqualFunctor (compare_functors.h:34-48).  For floating point it is
//   inf on either side -> a == b;  NaN -> false;  else fabs(double(a - b)) < 1e-8
// with a - b formed in the storage type.  (a == b) || |a-b| < 1e-8 is the same predicate:
// equal infinities give a - b = NaN, any NaN fails both tests.  For float, "(double)|d| <
// 1e-8" holds exactly when |d| <= 1e-8f, the largest float below 1e-8.
template <typename T, typename C>
__device__ __forceinline__ bool phi_equal(C a, C b) {
  if constexpr (std::numeric_limits<C>::is_integer) {
    return a == b;
  } else if constexpr (sizeof(C) == 8) {
    return (a == b) || (fabs(a - b) < 1e-8);
  } else {
    float d = a - b;
    if constexpr (sizeof(T) == 2) {
      // a - b is rounded to the 16-bit storage type before the test
      if constexpr (std::is_same<T, __nv_bfloat16>::value)
        d = __bfloat162float(__float2bfloat16_rn(d));
      else
        d = __half2float(__float2half_rn(d));
    }
    return (a == b) || (fabsf(d) <= 1e-8f);
  }
}

// attribute rounded to the storage type, as static_cast<T>(alpha) does in the reference
template <typename T>
__device__ __forceinline__ float attr_as(float v) {
  if constexpr (std::is_same<T, __nv_bfloat16>::value) return __bfloat162float(__float2bfloat16_rn(v));
  if constexpr (std::is_same<T, __half>::value) return __half2float(__float2half_rn(v));
  return v;
}

// 16-bit inputs: sigmoid / silu / tanh through the special-function unit's exp2 and reciprocal
// (relative error ~2^-21, against the 2^-9 the result is rounded to anyway) instead of expf and an
// IEEE division -- ~25 instructions per element, which made these functors issue-bound at twice
// the elements per byte (bf16 silu 4.2 TB/s where relu runs 6.4).  fp32 / fp64 keep the exact forms.
__device__ __forceinline__ float map_ex2_approx(float v) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
  return r;
}
__device__ __forceinline__ float map_rcp_approx(float v) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
  return r;
}
__device__ __forceinline__ float map_fast_sigmoid(float a) {
  return map_rcp_approx(1.f + map_ex2_approx(a * -1.4426950408889634f));
}

template <int kOp, typename T>
struct MapF {
  float p0, p1;

  // C: compute type of T (fp32 for 16-bit floats, int for narrow ints).  Returns C, or
  // 0/1 in an int for the BOOL-valued ops.
  template <typename C>
  __device__ __forceinline__ auto operator()(C a, C b, C c) const {
    constexpr bool kInt = std::numeric_limits<C>::is_integer;
    if constexpr (kOp == B200PHI_MAP_EQUAL) {
      return static_cast<int>(phi_equal<T, C>(a, b));
    } else if constexpr (kOp == B200PHI_MAP_NOT_EQUAL) {
      return static_cast<int>(!phi_equal<T, C>(a, b));
    } else if constexpr (kOp == B200PHI_MAP_LESS_THAN) {
      return static_cast<int>(a < b);
    } else if constexpr (kOp == B200PHI_MAP_LESS_EQUAL) {
      return static_cast<int>(a <= b);
    } else if constexpr (kOp == B200PHI_MAP_GREATER_THAN) {
      return static_cast<int>(a > b);
    } else if constexpr (kOp == B200PHI_MAP_GREATER_EQUAL) {
      return static_cast<int>(a >= b);
    } else if constexpr (kOp == B200PHI_MAP_LOGICAL_AND) {
      return static_cast<int>((a != C(0)) && (b != C(0)));
    } else if constexpr (kOp == B200PHI_MAP_LOGICAL_OR) {
      return static_cast<int>((a != C(0)) || (b != C(0)));
    } else if constexpr (kOp == B200PHI_MAP_LOGICAL_XOR) {
      return static_cast<int>((a != C(0)) != (b != C(0)));
    } else if constexpr (kOp == B200PHI_MAP_LOGICAL_NOT) {
      return static_cast<int>(!(a != C(0)));
    } else if constexpr (kOp == B200PHI_MAP_BITWISE_AND) {
      if constexpr (kInt) return static_cast<C>(a & b);
      else return a;
    } else if constexpr (kOp == B200PHI_MAP_BITWISE_OR) {
      if constexpr (kInt) return static_cast<C>(a | b);
      else return a;
    } else if constexpr (kOp == B200PHI_MAP_BITWISE_XOR) {
      if constexpr (kInt) return static_cast<C>(a ^ b);
      else return a;
    } else if constexpr (kOp == B200PHI_MAP_BITWISE_NOT) {
      if constexpr (std::is_same<T, Bool8>::value) return static_cast<C>(a == 0);
      else if constexpr (kInt) return static_cast<C>(~a);
      else return a;
    } else if constexpr (kOp >= B200PHI_MAP_LEFT_SHIFT_ARITHMETIC &&
                         kOp <= B200PHI_MAP_RIGHT_SHIFT_LOGIC) {
      if constexpr (kInt && !std::is_same<T, Bool8>::value) {
        constexpr int kBits = static_cast<int>(sizeof(T)) * 8;
        const bool in_range = b >= C(0) && b < C(kBits);
        if constexpr (kOp == B200PHI_MAP_LEFT_SHIFT_ARITHMETIC || kOp == B200PHI_MAP_LEFT_SHIFT_LOGIC) {
          // shifted in the unsigned image of C, truncated to T by the store
          using UC = typename std::make_unsigned<C>::type;
          return in_range ? static_cast<C>(static_cast<UC>(a) << static_cast<int>(b)) : C(0);
        } else if constexpr (kOp == B200PHI_MAP_RIGHT_SHIFT_ARITHMETIC) {
          if constexpr (std::is_unsigned<T>::value) return in_range ? static_cast<C>(a >> static_cast<int>(b)) : C(0);
          else return in_range ? static_cast<C>(a >> static_cast<int>(b)) : static_cast<C>(a < C(0) ? -1 : 0);
        } else {
          // logical: shift the unsigned image of the STORAGE type (logic_shift_func, :137-146)
          using UT = typename std::make_unsigned<T>::type;
          return in_range ? static_cast<C>(static_cast<UT>(static_cast<UT>(a) >> static_cast<int>(b))) : C(0);
        }
      } else {
        return a;
      }
    } else if constexpr (kOp == B200PHI_MAP_EQUAL_MUL) {
      return static_cast<C>(c * static_cast<C>(phi_equal<T, C>(a, b)));
    } else if constexpr (kOp == B200PHI_MAP_MAX_GRAD_X) {
      return static_cast<C>(c * static_cast<C>(a > b));
    } else if constexpr (kOp == B200PHI_MAP_MAX_GRAD_Y) {
      return static_cast<C>(c * static_cast<C>(a <= b));
    } else if constexpr (kOp == B200PHI_MAP_MIN_GRAD_X) {
      return static_cast<C>(c * static_cast<C>(a < b));
    } else if constexpr (kOp == B200PHI_MAP_MIN_GRAD_Y) {
      return static_cast<C>(c * static_cast<C>(a >= b));
    } else if constexpr (kOp == B200PHI_MAP_DIV_GRAD_Y) {
      if constexpr (kInt) return static_cast<C>(c == 0 ? 0 : -a * ((b / c) / c));
      else return static_cast<C>(-a * ((b / c) / c));
    } else if constexpr (kOp == B200PHI_MAP_EQUAL_T) {
      return static_cast<C>(phi_equal<T, C>(a, b));
    } else if constexpr (kOp == B200PHI_MAP_HEAVISIDE) {
      return a == C(0) ? b : static_cast<C>(a > C(0));
    } else if constexpr (kOp == B200PHI_MAP_REMAINDER) {  // RemainderFunctor, :540-592
      if constexpr (kInt) {
        if (b == C(0)) return C(0);
        C res = a % b;
        if ((res != C(0)) && ((b ^ res) < C(0))) res += b;
        return res;
      } else {
        C res = fmod(a, b);
        if ((res != C(0)) && ((res < C(0)) != (b < C(0)))) res += b;
        return res;
      }
    } else if constexpr (kOp == B200PHI_MAP_FLOOR_DIVIDE) {  // FloorDivideFunctor, :621-734
      if constexpr (kInt) {
        if (b == C(0)) return C(0);
        if ((a < C(0)) != (b < C(0))) {
          const C quot = a / b, rem = a % b;
          return rem ? static_cast<C>(quot - 1) : quot;
        }
        return static_cast<C>(a / b);
      } else {
        if (b == C(0)) return static_cast<C>(a / b);  // the IEEE result
        const C mod = fmod(a, b);
        C div = (a - mod) / b;
        if ((mod != C(0)) && ((b < C(0)) != (mod < C(0)))) div -= C(1);
        C fd;
        if (div != C(0)) {
          fd = floor(div);
          if (div - fd > C(0.5)) fd += C(1);
        } else {
          fd = copysign(C(0), a / b);
        }
        return fd;
      }
    } else if constexpr (kOp == B200PHI_MAP_POW) {  // ElementwisePowFunctor, :844-889 (device branch)
      if constexpr (kInt) return static_cast<C>(llrint(pow(static_cast<double>(a), static_cast<double>(b))));
      else return pow(a, b);
    } else if constexpr (kOp == B200PHI_MAP_COPYSIGN) {  // copysign_func, :903-921
      if constexpr (kInt) return static_cast<C>(copysign(static_cast<double>(a), static_cast<double>(b)));
      else return copysign(a, b);
    } else if constexpr (kOp == B200PHI_MAP_FMAX || kOp == B200PHI_MAP_FMIN) {  // :190-290
      constexpr bool kMx = kOp == B200PHI_MAP_FMAX;
      if constexpr (std::is_same<T, int32_t>::value) {
        const float fa = static_cast<float>(a), fb = static_cast<float>(b);
        return static_cast<C>(lrintf(kMx ? fmaxf(fa, fb) : fminf(fa, fb)));
      } else if constexpr (kInt) {
        const double fa = static_cast<double>(a), fb = static_cast<double>(b);
        return static_cast<C>(llrint(kMx ? fmax(fa, fb) : fmin(fa, fb)));
      } else {
        return kMx ? fmax(a, b) : fmin(a, b);
      }
    } else if constexpr (kOp == B200PHI_MAP_ABS) {
      return static_cast<C>(a < C(0) ? -a : a);
    } else if constexpr (kOp == B200PHI_MAP_RELU) {
      return static_cast<C>(a > C(0) ? a : C(0));
    } else if constexpr (kInt) {
      return a;  // floating-point only below; refused on the host for integer dtypes
    } else {
      const C one = C(1), zero = C(0);
      if constexpr (kOp == B200PHI_MAP_SUB_EXP) return exp(a - b);
      else if constexpr (kOp == B200PHI_MAP_EXP) return exp(a);
      else if constexpr (kOp == B200PHI_MAP_LOG) return log(a);
      else if constexpr (kOp == B200PHI_MAP_SQRT) return sqrt(a);
      else if constexpr (kOp == B200PHI_MAP_RSQRT) return rsqrt(a);
      else if constexpr (kOp == B200PHI_MAP_SQUARE) return a * a;
      else if constexpr (kOp == B200PHI_MAP_RECIPROCAL) return one / a;
      else if constexpr (kOp == B200PHI_MAP_SIN) return sin(a);
      else if constexpr (kOp == B200PHI_MAP_COS) return cos(a);
      else if constexpr (kOp == B200PHI_MAP_TANH) {
        if constexpr (sizeof(T) == 2) {  // (see map_fast_sigmoid)
          const float x = fabsf(a);
          const float t = map_ex2_approx(x * -2.8853900817779268f);  // e^(-2|a|)
          const float big = (1.f - t) * map_rcp_approx(1.f + t);
          const float small = x * (1.f - x * x * 0.3333333333f);      // |a| < 0.04: no cancellation
          return copysignf(x < 0.04f ? small : big, a);
        } else {
          return tanh(a);
        }
      } else if constexpr (kOp == B200PHI_MAP_SIGMOID) {
        if constexpr (sizeof(T) == 2) return map_fast_sigmoid(a);
        else return one / (one + exp(-a));
      } else if constexpr (kOp == B200PHI_MAP_SILU) {
        if constexpr (sizeof(T) == 2) return a * map_fast_sigmoid(a);
        else return a / (one + exp(-a));
      }
      else if constexpr (kOp == B200PHI_MAP_LEAKY_RELU)
        return a > zero ? a : static_cast<C>(attr_as<T>(p0)) * a;
      else if constexpr (kOp == B200PHI_MAP_ELU)
        return a > zero ? a : static_cast<C>(p0) * (exp(a) - one);
      else if constexpr (kOp == B200PHI_MAP_SOFTPLUS) {
        const C beta = static_cast<C>(p0), xb = a * beta;
        return xb > static_cast<C>(p1) ? a : log(one + exp(xb)) / beta;
      } else if constexpr (kOp == B200PHI_MAP_FLOOR) return floor(a);
      else if constexpr (kOp == B200PHI_MAP_CEIL) return ceil(a);
      else if constexpr (kOp == B200PHI_MAP_EXPM1) return expm1(a);
      else if constexpr (kOp == B200PHI_MAP_LOG1P) return log(one + a);
      else if constexpr (kOp == B200PHI_MAP_LOG2) return log2(a);
      else if constexpr (kOp == B200PHI_MAP_LOG10) return log10(a);
      else if constexpr (kOp == B200PHI_MAP_RELU6) {
        const C t = static_cast<C>(attr_as<T>(p0));
        return a <= zero ? zero : (a < t ? a : t);
      } else if constexpr (kOp == B200PHI_MAP_HARD_SIGMOID) {
        const C t = a * static_cast<C>(attr_as<T>(p0)) + static_cast<C>(attr_as<T>(p1));
        const C hi = t > zero ? t : zero;
        return hi < one ? hi : one;
      } else if constexpr (kOp == B200PHI_MAP_SWISH) {
        if constexpr (sizeof(T) == 2) return a * map_fast_sigmoid(static_cast<C>(p0) * a);
        else return a / (one + exp(-static_cast<C>(p0) * a));
      }
      else if constexpr (kOp == B200PHI_MAP_MISH) {
        const C sp = a > static_cast<C>(p0) ? a : log(one + exp(a));
        return a * tanh(sp);
      } else if constexpr (kOp == B200PHI_MAP_SOFTSIGN) return a / (one + (a > -a ? a : -a));
      else if constexpr (kOp == B200PHI_MAP_LOG_SIGMOID) {
        const C t = a > zero ? zero : -a;
        return -t - log(exp(-t) + exp(-a - t));
      } else if constexpr (kOp == B200PHI_MAP_POW_ABS)
        return pow(a < zero ? -a : a, static_cast<C>(p0));
      else return a;
    }
  }
};

}  // namespace b200phi
