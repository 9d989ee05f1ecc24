// =============================================================================
// oracle/ref_shim/float_grad.h -- TEST INFRASTRUCTURE ONLY.
//
// Functional stand-in for the MISSING third-party header `float_grad.h` of the reference
// (module rogerhh/auto-jvp-example @ 06f0e5508bdf2ed3548c34c0f3a943027736cb83, a private
// submodule absent from the reference archive -- SURVEY.md F1).  It restates the published
// semantics of that library -- scalar forward-mode dual numbers: a value and a tangent carried
// through + - * / sqrt exp ceil, comparisons and branches decided on the value -- behind the
// names the reference's own headers use (call sites: DGR/cuda_rasterizer/forward_impl.h,
// float_grad_{vec3,vec4,mat3,mat4,helper_math}.h, helper_math.h:70-84, auxiliary.h:95-259,
// rasterizer_jvp_impl.h, rasterizer.h:144-185, rasterize_points.cu:244-259), so that the
// reference's UNMODIFIED JVP kernels (preprocessCUDAJvp, renderCUDAJvpSplit, forwardJvp) can be
// compiled from where they lie into oracle/_ref/ and used as the GPU-side parity oracle for the
// forward-mode half of J^T J v.  The shim itself is pinned on the reference's golden vectors:
// preprocessCUDAJvp built with it must reproduce DGR/tests/render_data/*_grad.csv
// (tests/test_gpu_vs_reference.py::test_ref_jvp_shim_reproduces_fixture).
//
// Nothing of the product includes this file.
// =============================================================================
#pragma once
#include <cuda_runtime.h>
#include <cmath>
#include <initializer_list>
#include <type_traits>
#include <utility>

#define FG_HD inline __host__ __device__

template <typename T> struct FloatGradBase;
template <typename T> struct FloatGradRefBase;
template <typename T> struct FloatGrad;
template <typename T> struct FloatGradRef;
template <typename T> struct FloatGradArray;

template <typename T> struct always_false : std::false_type {};

// ---- classification --------------------------------------------------------------------------
namespace fg_detail {
template <typename T> struct kind { static constexpr int value = 0; };            // 0 plain, 1 value, 2 ref, 3 array
template <typename T> struct kind<FloatGrad<T>> { static constexpr int value = 1; };
template <typename T> struct kind<FloatGradRef<T>> { static constexpr int value = 2; };
template <typename T> struct kind<FloatGradArray<T>> { static constexpr int value = 3; };
template <typename T> using bare = std::remove_cv_t<std::remove_reference_t<T>>;
template <typename T> constexpr int kind_v = kind<bare<T>>::value;
template <typename T> constexpr bool is_dual_v = kind_v<T> == 1 || kind_v<T> == 2;   // carries ONE value + tangent
template <typename V> constexpr bool is_cuda_vec_v =
    std::is_same<V, float2>::value || std::is_same<V, float3>::value || std::is_same<V, float4>::value;

template <typename V> FG_HD V zero() {
    if constexpr (std::is_arithmetic<V>::value) return V(0);
    else if constexpr (is_cuda_vec_v<V>) return V{};
    else return V(0.0f);                       // glm vectors / matrices
}
// build a V from components (CUDA vector types have no constructors; glm types do)
template <typename V, typename... C> FG_HD V build(C... c) {
    if constexpr (std::is_same<V, float2>::value) return make_float2((float)c...);
    else if constexpr (std::is_same<V, float3>::value) return make_float3((float)c...);
    else if constexpr (std::is_same<V, float4>::value) return make_float4((float)c...);
    else return V(c...);
}
// a tangent brought to the shape of the result (scalar tangent of `vector + scalar` is the same for every component)
template <typename R, typename X> FG_HD R as(const X& x) {
    if constexpr (std::is_same<R, bare<X>>::value) return x;
    else if constexpr (std::is_same<R, float2>::value) return make_float2((float)x, (float)x);
    else if constexpr (std::is_same<R, float3>::value) return make_float3((float)x, (float)x, (float)x);
    else if constexpr (std::is_same<R, float4>::value) return make_float4((float)x, (float)x, (float)x, (float)x);
    else return R(x);
}
}  // namespace fg_detail

// true for everything that carries tangents, arrays included (the reference selects dual result types with
// is_float_grad<decltype(array argument)>, forward_impl.h:39-41,117-141,280-297)
template <typename T> struct is_float_grad : std::integral_constant<bool, fg_detail::kind_v<T> != 0> {};

// ---- value / tangent access --------------------------------------------------------------------
template <typename T> FG_HD decltype(auto) get_data(const T& v) {
    if constexpr (fg_detail::is_dual_v<T>) return v.data();
    else return (v);
}
template <typename T> FG_HD decltype(auto) get_data(T& v) {
    if constexpr (fg_detail::is_dual_v<T>) return v.data();
    else return (v);
}
// primary template; the reference specialises it for glm types (float_grad_vec3.h:8-17, ...)
template <typename T> FG_HD decltype(auto) get_grad(const T& v) {
    if constexpr (fg_detail::is_dual_v<T>) return v.grad();
    else return fg_detail::zero<std::remove_cv_t<T>>();
}
template <typename T> FG_HD auto get_data_ptr(const T& a) {
    if constexpr (fg_detail::kind_v<T> == 3) return a.data_ptr();
    else return a;
}
template <typename T> FG_HD auto get_grad_ptr(const T& a) {
    if constexpr (fg_detail::kind_v<T> == 3) return a.grad_ptr();
    else return a;
}

namespace fg_detail {
template <typename T> using data_t = bare<decltype(get_data(std::declval<const bare<T>&>()))>;
template <typename T> constexpr bool is_scalar_like_v = std::is_arithmetic<data_t<T>>::value && kind_v<T> != 3;
// tangent of one scalar component argument, as float-compatible value
template <typename A> FG_HD auto comp_grad(const A& a) {
    if constexpr (is_dual_v<A>) return a.grad();
    else return 0.0f;
}
template <typename V, typename A> FG_HD V grad_as(const A& a) {
    if constexpr (is_dual_v<A>) return V(a.grad());
    else return zero<V>();
}
}  // namespace fg_detail

template <typename T> using is_float_type = std::is_same<fg_detail::data_t<T>, float>;
template <typename T> using is_float2_type = std::is_same<fg_detail::data_t<T>, float2>;
template <typename T> using is_float3_type = std::is_same<fg_detail::data_t<T>, float3>;
template <typename T> using is_float4_type = std::is_same<fg_detail::data_t<T>, float4>;

// ---- storage ---------------------------------------------------------------------------------
// Owning dual: value + tangent.  Constructor forms used by the reference:
//   (x)            value, zero tangent (x plain) / copy of both (x dual)
//   (value, tangent)   when both are plain and already of the stored type (or the stored type is a scalar)
//   (c0, c1, ...)  component-wise from plain or dual scalars / columns
template <typename T> struct FloatGradBase {
    using V = std::remove_const_t<T>;
    V data_, grad_;
    FG_HD FloatGradBase() : data_(fg_detail::zero<V>()), grad_(fg_detail::zero<V>()) {}
    template <typename A> FG_HD FloatGradBase(const A& a) : data_(V(get_data(a))), grad_(fg_detail::grad_as<V>(a)) {}
    template <typename A, typename B> FG_HD FloatGradBase(const A& a, const B& b) {
        constexpr bool pair = !fg_detail::is_dual_v<A> && !fg_detail::is_dual_v<B> &&
                              (std::is_arithmetic<V>::value ||
                               (std::is_same<fg_detail::bare<A>, V>::value && std::is_same<fg_detail::bare<B>, V>::value));
        if constexpr (pair) { data_ = V(a); grad_ = V(b); }
        else {
            data_ = fg_detail::build<V>(get_data(a), get_data(b));
            grad_ = fg_detail::build<V>(fg_detail::comp_grad(a), fg_detail::comp_grad(b));
        }
    }
    template <typename A, typename B, typename C, typename... R> FG_HD FloatGradBase(const A& a, const B& b, const C& c, const R&... r)
        : data_(fg_detail::build<V>(get_data(a), get_data(b), get_data(c), get_data(r)...)),
          grad_(fg_detail::build<V>(fg_detail::comp_grad(a), fg_detail::comp_grad(b), fg_detail::comp_grad(c),
                                    fg_detail::comp_grad(r)...)) {}
    FG_HD T& data() { return data_; }
    FG_HD const V& data() const { return data_; }
    FG_HD T& grad() { return grad_; }
    FG_HD const V& grad() const { return grad_; }
    template <typename O> FG_HD FloatGradBase& operator=(const O& o) {
        data_ = V(get_data(o));
        grad_ = fg_detail::grad_as<V>(o);
        return *this;
    }
};

// Non-owning dual: pointers to a value and to its tangent; assignment writes through.
template <typename T> struct FloatGradRefBase {
    using V = std::remove_const_t<T>;
    T* data_ptr_;
    T* grad_ptr_;
    FG_HD FloatGradRefBase(T* d, T* g) : data_ptr_(d), grad_ptr_(g) {}
    template <typename U, typename = std::enable_if_t<std::is_same<std::remove_const_t<U>, V>::value>>
    FG_HD FloatGradRefBase(const FloatGradRefBase<U>& o) : data_ptr_(o.data_ptr_), grad_ptr_(o.grad_ptr_) {}
    template <typename U, typename = std::enable_if_t<std::is_same<std::remove_const_t<U>, V>::value>>
    FG_HD FloatGradRefBase(FloatGradBase<U>& o) : data_ptr_(&o.data_), grad_ptr_(&o.grad_) {}
    FG_HD T& data() const { return *data_ptr_; }
    FG_HD T& grad() const { return *grad_ptr_; }
    FloatGradRefBase(const FloatGradRefBase&) = default;
    FG_HD FloatGradRefBase& operator=(const FloatGradRefBase& o) {     // a reference is assigned THROUGH, never re-seated
        *data_ptr_ = *o.data_ptr_;
        *grad_ptr_ = *o.grad_ptr_;
        return *this;
    }
    template <typename O> FG_HD const FloatGradRefBase& operator=(const O& o) const {
        *data_ptr_ = V(get_data(o));
        *grad_ptr_ = fg_detail::grad_as<V>(o);
        return *this;
    }
};

// ---- primary templates: scalars ----------------------------------------------------------------
template <typename T> struct FloatGrad : FloatGradBase<T> {
    using Base = FloatGradBase<T>;
    FG_HD FloatGrad() : Base() {}
    template <typename A> FG_HD FloatGrad(const A& a) : Base(a) {}
    template <typename A, typename B> FG_HD FloatGrad(const A& a, const B& b) : Base(a, b) {}
    template <typename O> FG_HD FloatGrad& operator=(const O& o) { Base::operator=(o); return *this; }
};
template <typename T> struct FloatGradRef : FloatGradRefBase<T> {
    using Base = FloatGradRefBase<T>;
    FG_HD FloatGradRef(T* d, T* g) : Base(d, g) {}
    FG_HD FloatGradRef(const FloatGradRef& o) : Base(o.data_ptr_, o.grad_ptr_) {}
    template <typename U> FG_HD FloatGradRef(const FloatGradRef<U>& o) : Base(o.data_ptr_, o.grad_ptr_) {}
    template <typename U> FG_HD FloatGradRef(FloatGrad<U>& o) : Base(&o.data_, &o.grad_) {}
    FG_HD const FloatGradRef& operator=(const FloatGradRef& o) const { Base::operator=(o); return *this; }   // write through
    template <typename O> FG_HD const FloatGradRef& operator=(const O& o) const { Base::operator=(o); return *this; }
};

// ---- CUDA vector types: components are references into the owner ---------------------------------
#define FG_VEC_COMMON(NAME, BASE, VT, CT)                                                             \
    FG_HD FloatGradRef<CT> operator[](int i) const {                                                  \
        return FloatGradRef<CT>(&(&this->data().x)[i], &(&this->grad().x)[i]);                        \
    }

#define FG_DEFINE_VEC(VT, INIT, ...)                                                                   \
    template <> struct FloatGrad<VT> : FloatGradBase<VT> {                                             \
        using Base = FloatGradBase<VT>;                                                                \
        FG_HD FloatGrad() : Base(), INIT {}                                                            \
        FG_HD FloatGrad(const FloatGrad& o) : Base(o.data_, o.grad_), INIT {}                          \
        template <typename... A> FG_HD FloatGrad(const A&... a) : Base(a...), INIT {}                  \
        FG_HD FloatGrad& operator=(const FloatGrad& o) { data_ = o.data_; grad_ = o.grad_; return *this; } \
        template <typename O> FG_HD FloatGrad& operator=(const O& o) { Base::operator=(o); return *this; } \
        FG_HD FloatGradRef<float> operator[](int i) { return FloatGradRef<float>(&(&data_.x)[i], &(&grad_.x)[i]); } \
        FloatGradRef<float> __VA_ARGS__;                                                               \
    };                                                                                                 \
    template <> struct FloatGradRef<VT> : FloatGradRefBase<VT> {                                       \
        using Base = FloatGradRefBase<VT>;                                                             \
        FG_HD FloatGradRef(VT* d, VT* g) : Base(d, g), INIT {}                                         \
        FG_HD FloatGradRef(const FloatGradRef& o) : Base(o.data_ptr_, o.grad_ptr_), INIT {}            \
        FG_HD FloatGradRef(FloatGrad<VT>& o) : Base(&o.data_, &o.grad_), INIT {}                       \
        FG_HD const FloatGradRef& operator=(const FloatGradRef& o) const { Base::operator=(o); return *this; } \
        template <typename O> FG_HD const FloatGradRef& operator=(const O& o) const { Base::operator=(o); return *this; } \
        FG_HD FloatGradRef<float> operator[](int i) const { return FloatGradRef<float>(&(&data().x)[i], &(&grad().x)[i]); } \
        FloatGradRef<float> __VA_ARGS__;                                                               \
    };                                                                                                 \
    template <> struct FloatGradRef<const VT> : FloatGradRefBase<const VT> {                           \
        using Base = FloatGradRefBase<const VT>;                                                       \
        FG_HD FloatGradRef(const VT* d, const VT* g) : Base(d, g), INIT {}                             \
        FG_HD FloatGradRef(const FloatGradRef& o) : Base(o.data_ptr_, o.grad_ptr_), INIT {}            \
        FG_HD FloatGradRef(const FloatGradRef<VT>& o) : Base(o.data_ptr_, o.grad_ptr_), INIT {}        \
        FG_HD FloatGradRef<const float> operator[](int i) const { return FloatGradRef<const float>(&(&data().x)[i], &(&grad().x)[i]); } \
        FloatGradRef<const float> __VA_ARGS__;                                                         \
    };

#define FG_COMMA ,
FG_DEFINE_VEC(float2, x(&this->data().x FG_COMMA &this->grad().x) FG_COMMA y(&this->data().y FG_COMMA &this->grad().y), x, y)
FG_DEFINE_VEC(float3, x(&this->data().x FG_COMMA &this->grad().x) FG_COMMA y(&this->data().y FG_COMMA &this->grad().y) FG_COMMA
                          z(&this->data().z FG_COMMA &this->grad().z), x, y, z)
FG_DEFINE_VEC(float4, x(&this->data().x FG_COMMA &this->grad().x) FG_COMMA y(&this->data().y FG_COMMA &this->grad().y) FG_COMMA
                          z(&this->data().z FG_COMMA &this->grad().z) FG_COMMA w(&this->data().w FG_COMMA &this->grad().w), x, y, z, w)
#undef FG_DEFINE_VEC
#undef FG_VEC_COMMON

// ---- arrays: a pair of pointers (values, tangents); trivially copyable so it can be a kernel argument ----------
template <typename T> struct FloatGradArray {
    T* data_arr_ = nullptr;
    T* grad_arr_ = nullptr;
    FloatGradArray() = default;
    FG_HD FloatGradArray(T* d, T* g) : data_arr_(d), grad_arr_(g) {}
    FG_HD FloatGradArray(std::nullptr_t) : data_arr_(nullptr), grad_arr_(nullptr) {}
    template <typename U, typename = std::enable_if_t<std::is_same<const U, T>::value>>
    FG_HD FloatGradArray(const FloatGradArray<U>& o) : data_arr_(o.data_arr_), grad_arr_(o.grad_arr_) {}
    FG_HD T* data_ptr() const { return data_arr_; }
    FG_HD T* grad_ptr() const { return grad_arr_; }
    FG_HD FloatGradRef<T> operator[](size_t i) const { return FloatGradRef<T>(data_arr_ + i, grad_arr_ + i); }
    FG_HD FloatGradRef<T> operator*() const { return FloatGradRef<T>(data_arr_, grad_arr_); }
    FG_HD FloatGradArray operator+(ptrdiff_t n) const { return FloatGradArray(data_arr_ + n, grad_arr_ + n); }
    FG_HD bool operator==(std::nullptr_t) const { return data_arr_ == nullptr; }
    FG_HD bool operator!=(std::nullptr_t) const { return data_arr_ != nullptr; }
};
template <typename T> using SharedFloatGradArray = FloatGradArray<T>;

// cast<glm::vec3>(array of float) -> array of glm::vec3 over the same memory (constness kept)
template <typename To, typename From> FG_HD auto cast(const FloatGradArray<From>& a) {
    using R = std::conditional_t<std::is_const<From>::value, const To, To>;
    return FloatGradArray<R>(reinterpret_cast<R*>(a.data_arr_), reinterpret_cast<R*>(a.grad_arr_));
}
template <typename To, typename From> FG_HD auto cast(From* p) {
    using R = std::conditional_t<std::is_const<From>::value, const To, To>;
    return reinterpret_cast<R*>(p);
}

// ---- arithmetic: any operator with at least one dual operand ---------------------------------------------------
#define FG_ENABLE_BIN(A, B) std::enable_if_t<(fg_detail::is_dual_v<A> || fg_detail::is_dual_v<B>) && \
                                             fg_detail::kind_v<A> != 3 && fg_detail::kind_v<B> != 3, int> = 0

template <typename A, typename B, FG_ENABLE_BIN(A, B)> FG_HD auto operator+(const A& a, const B& b) {
    using R = fg_detail::bare<decltype(get_data(a) + get_data(b))>;
    if constexpr (fg_detail::is_dual_v<A> && fg_detail::is_dual_v<B>)
        return FloatGrad<R>(R(get_data(a) + get_data(b)), R(a.grad() + b.grad()));
    else if constexpr (fg_detail::is_dual_v<A>) return FloatGrad<R>(R(get_data(a) + b), fg_detail::as<R>(a.grad()));
    else return FloatGrad<R>(R(a + get_data(b)), fg_detail::as<R>(b.grad()));
}
template <typename A, typename B, FG_ENABLE_BIN(A, B)> FG_HD auto operator-(const A& a, const B& b) {
    using R = fg_detail::bare<decltype(get_data(a) - get_data(b))>;
    if constexpr (fg_detail::is_dual_v<A> && fg_detail::is_dual_v<B>)
        return FloatGrad<R>(R(get_data(a) - get_data(b)), R(a.grad() - b.grad()));
    else if constexpr (fg_detail::is_dual_v<A>) return FloatGrad<R>(R(get_data(a) - b), fg_detail::as<R>(a.grad()));
    else return FloatGrad<R>(R(a - get_data(b)), fg_detail::as<R>(-b.grad()));
}
template <typename A, typename B, FG_ENABLE_BIN(A, B)> FG_HD auto operator*(const A& a, const B& b) {
    using R = fg_detail::bare<decltype(get_data(a) * get_data(b))>;
    if constexpr (fg_detail::is_dual_v<A> && fg_detail::is_dual_v<B>)
        return FloatGrad<R>(R(get_data(a) * get_data(b)), R(a.grad() * get_data(b) + get_data(a) * b.grad()));
    else if constexpr (fg_detail::is_dual_v<A>) return FloatGrad<R>(R(get_data(a) * b), R(a.grad() * b));
    else return FloatGrad<R>(R(a * get_data(b)), R(a * b.grad()));
}
template <typename A, typename B, FG_ENABLE_BIN(A, B)> FG_HD auto operator/(const A& a, const B& b) {
    using R = fg_detail::bare<decltype(get_data(a) / get_data(b))>;
    if constexpr (fg_detail::is_dual_v<A> && fg_detail::is_dual_v<B>) {
        const auto q = get_data(a) / get_data(b);
        return FloatGrad<R>(R(q), R((a.grad() - q * b.grad()) / get_data(b)));
    } else if constexpr (fg_detail::is_dual_v<A>) return FloatGrad<R>(R(get_data(a) / b), R(a.grad() / b));
    else {
        const auto q = a / get_data(b);
        return FloatGrad<R>(R(q), R(-(q * b.grad()) / get_data(b)));
    }
}
template <typename A, std::enable_if_t<fg_detail::is_dual_v<A> && fg_detail::is_scalar_like_v<A>, int> = 0>
FG_HD auto operator-(const A& a) {
    using R = fg_detail::data_t<A>;
    return FloatGrad<R>(R(-a.data()), R(-a.grad()));
}
template <typename A, std::enable_if_t<fg_detail::is_dual_v<A> && fg_detail::is_scalar_like_v<A>, int> = 0>
FG_HD auto operator+(const A& a) {
    using R = fg_detail::data_t<A>;
    return FloatGrad<R>(R(a.data()), R(a.grad()));
}
#define FG_COMPOUND(OP)                                                                                           \
    template <typename A, typename B, std::enable_if_t<fg_detail::is_dual_v<A>, int> = 0>                          \
    FG_HD decltype(auto) operator OP##=(A&& a, const B& b) { a = a OP b; return std::forward<A>(a); }
FG_COMPOUND(+) FG_COMPOUND(-) FG_COMPOUND(*) FG_COMPOUND(/)
#undef FG_COMPOUND

// comparisons: decided on the value (branches of the reference's dual kernels follow the primal)
#define FG_COMPARE(OP)                                                                                            \
    template <typename A, typename B, FG_ENABLE_BIN(A, B)> FG_HD bool operator OP(const A& a, const B& b) {       \
        return get_data(a) OP get_data(b);                                                                        \
    }
FG_COMPARE(<) FG_COMPARE(>) FG_COMPARE(<=) FG_COMPARE(>=) FG_COMPARE(==) FG_COMPARE(!=)
#undef FG_COMPARE

// ---- elementary functions on scalar duals ------------------------------------------------------------------------
#define FG_ENABLE_SCALAR(A) std::enable_if_t<fg_detail::is_dual_v<A> && fg_detail::is_scalar_like_v<A>, int> = 0
template <typename A, FG_ENABLE_SCALAR(A)> FG_HD FloatGrad<float> sqrtf(const A& a) {
    const float s = ::sqrtf((float)a.data());
    return FloatGrad<float>(s, (float)a.grad() / (2.0f * s));
}
template <typename A, FG_ENABLE_SCALAR(A)> FG_HD FloatGrad<float> sqrt(const A& a) { return sqrtf(a); }
template <typename A, FG_ENABLE_SCALAR(A)> FG_HD FloatGrad<float> expf(const A& a) {
    const float e = ::expf((float)a.data());
    return FloatGrad<float>(e, e * (float)a.grad());
}
template <typename A, FG_ENABLE_SCALAR(A)> FG_HD FloatGrad<float> exp(const A& a) { return expf(a); }
template <typename A, FG_ENABLE_SCALAR(A)> FG_HD FloatGrad<float> ceil(const A& a) {
    return FloatGrad<float>(::ceilf((float)a.data()), 0.0f);
}
template <typename A, FG_ENABLE_SCALAR(A)> FG_HD FloatGrad<float> ceilf(const A& a) { return ceil(a); }
template <typename A, FG_ENABLE_SCALAR(A)> FG_HD FloatGrad<float> floorf(const A& a) {
    return FloatGrad<float>(::floorf((float)a.data()), 0.0f);
}
template <typename A, FG_ENABLE_SCALAR(A)> FG_HD FloatGrad<float> fabs(const A& a) {
    return a.data() < 0 ? FloatGrad<float>(-(float)a.data(), -(float)a.grad()) : FloatGrad<float>((float)a.data(), (float)a.grad());
}
template <typename A, FG_ENABLE_SCALAR(A)> FG_HD FloatGrad<float> fabsf(const A& a) { return fabs(a); }
// max / min pick the operand (value and tangent) with the larger / smaller value
template <typename A, typename B, std::enable_if_t<(fg_detail::is_dual_v<A> || fg_detail::is_dual_v<B>) &&
                                                   fg_detail::is_scalar_like_v<A> && fg_detail::is_scalar_like_v<B>, int> = 0>
FG_HD FloatGrad<float> max(const A& a, const B& b) {
    return get_data(a) > get_data(b) ? FloatGrad<float>(a) : FloatGrad<float>(b);
}
template <typename A, typename B, std::enable_if_t<(fg_detail::is_dual_v<A> || fg_detail::is_dual_v<B>) &&
                                                   fg_detail::is_scalar_like_v<A> && fg_detail::is_scalar_like_v<B>, int> = 0>
FG_HD FloatGrad<float> min(const A& a, const B& b) {
    return get_data(a) < get_data(b) ? FloatGrad<float>(a) : FloatGrad<float>(b);
}

// ---- make_floatN from scalar components, at least one of them dual ----------------------------------------------------
#define FG_ANY_DUAL(...) (fg_detail::is_dual_v<__VA_ARGS__>)
template <typename A, typename B,
          std::enable_if_t<(fg_detail::is_dual_v<A> || fg_detail::is_dual_v<B>) && fg_detail::is_scalar_like_v<A> &&
                           fg_detail::is_scalar_like_v<B>, int> = 0>
FG_HD FloatGrad<float2> make_float2(const A& a, const B& b) {
    return FloatGrad<float2>(::make_float2((float)get_data(a), (float)get_data(b)),
                             ::make_float2((float)fg_detail::comp_grad(a), (float)fg_detail::comp_grad(b)));
}
template <typename A, typename B, typename C,
          std::enable_if_t<(fg_detail::is_dual_v<A> || fg_detail::is_dual_v<B> || fg_detail::is_dual_v<C>) &&
                           fg_detail::is_scalar_like_v<A> && fg_detail::is_scalar_like_v<B> && fg_detail::is_scalar_like_v<C>, int> = 0>
FG_HD FloatGrad<float3> make_float3(const A& a, const B& b, const C& c) {
    return FloatGrad<float3>(::make_float3((float)get_data(a), (float)get_data(b), (float)get_data(c)),
                             ::make_float3((float)fg_detail::comp_grad(a), (float)fg_detail::comp_grad(b), (float)fg_detail::comp_grad(c)));
}
template <typename A, typename B, typename C, typename D,
          std::enable_if_t<(fg_detail::is_dual_v<A> || fg_detail::is_dual_v<B> || fg_detail::is_dual_v<C> || fg_detail::is_dual_v<D>) &&
                           fg_detail::is_scalar_like_v<A> && fg_detail::is_scalar_like_v<B> && fg_detail::is_scalar_like_v<C> &&
                           fg_detail::is_scalar_like_v<D>, int> = 0>
FG_HD FloatGrad<float4> make_float4(const A& a, const B& b, const C& c, const D& d) {
    return FloatGrad<float4>(::make_float4((float)get_data(a), (float)get_data(b), (float)get_data(c), (float)get_data(d)),
                             ::make_float4((float)fg_detail::comp_grad(a), (float)fg_detail::comp_grad(b),
                                           (float)fg_detail::comp_grad(c), (float)fg_detail::comp_grad(d)));
}
