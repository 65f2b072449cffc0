// TEST INFRASTRUCTURE (oracle): container typedefs over thrust vectors + strided_range.
#pragma once
#include <mcbooster/GTypes.h>
#include <mcbooster/Vector4R.h>
#include <thrust/device_vector.h>
#include <thrust/host_vector.h>
#include <thrust/functional.h>
#include <thrust/copy.h>
#include <thrust/count.h>
#include <thrust/extrema.h>
#include <thrust/random.h>
#include <thrust/reduce.h>
#include <thrust/remove.h>
#include <thrust/sequence.h>
#include <thrust/transform.h>
#include <thrust/transform_reduce.h>
#include <thrust/tuple.h>
#include <thrust/iterator/zip_iterator.h>
#include <thrust/iterator/constant_iterator.h>
#include <thrust/iterator/counting_iterator.h>
#include <thrust/iterator/permutation_iterator.h>
#include <thrust/iterator/transform_iterator.h>
#include <vector>
namespace mcbooster {
template <typename T> using mc_device_vector = thrust::device_vector<T>;
template <typename T> using mc_host_vector   = thrust::host_vector<T>;
typedef mc_device_vector<GReal_t> RealVector_d;
typedef mc_host_vector<GReal_t> RealVector_h;
typedef mc_device_vector<GBool_t> BoolVector_d;
typedef mc_host_vector<GBool_t> BoolVector_h;
typedef mc_device_vector<Vector4R> Particles_d;
typedef mc_host_vector<Vector4R> Particles_h;
typedef std::vector<Particles_d *> ParticlesSet_d;
typedef std::vector<Particles_h *> ParticlesSet_h;
typedef std::vector<RealVector_d *> VariableSet_d;
typedef std::vector<RealVector_h *> VariableSet_h;

template <typename Iterator> class strided_range {
  public:
    typedef typename thrust::iterator_difference<Iterator>::type difference_type;
    struct stride_functor {
        difference_type stride;
        stride_functor(difference_type s) : stride(s) {}
        __host__ __device__ difference_type operator()(const difference_type &i) const { return stride * i; }
    };
    typedef thrust::counting_iterator<difference_type> CountingIterator;
    typedef thrust::transform_iterator<stride_functor, CountingIterator> TransformIterator;
    typedef thrust::permutation_iterator<Iterator, TransformIterator> PermutationIterator;
    typedef PermutationIterator iterator;
    strided_range(Iterator first, Iterator last, difference_type stride) : first(first), last(last), stride(stride) {}
    iterator begin() const {
        return PermutationIterator(first, TransformIterator(CountingIterator(0), stride_functor(stride)));
    }
    iterator end() const { return begin() + ((last - first) + (stride - 1)) / stride; }
  protected:
    Iterator first, last;
    difference_type stride;
};
} // namespace mcbooster
