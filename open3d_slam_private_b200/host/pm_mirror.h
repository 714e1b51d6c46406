// pm_mirror.h — a MINIMAL stand-in for the parts of libpointmatcher's PointMatcher<float> interface that the B200
// shims (b200_pm_modules.hpp) touch. It exists only because this build image has no Eigen/Boost/libnabo, so
// libpointmatcher's own PointMatcher.h cannot be included here. The shims are templates over a `PM` type: in the
// reference tree they are instantiated with the real `PointMatcher<float>` (whose Eigen matrices offer the same
// data()/rows()/cols()/operator() used below); in this repo they are instantiated with `pmmirror::PM` so that they
// are compiled and exercised (host/selftest.cpp).
//
// Mirrors (names, members and meaning) of:  libpointmatcher/pointmatcher/PointMatcher.h
//   Matrix/IntMatrix (Eigen column-major)         :170-190      DataPoints (features, descriptors, labels) :222-403
//   Matches (dists, ids, InvalidId, InvalidDist)  :416-436      OutlierWeights                               :442
//   Matcher {visitCounter, init, findClosests}    :518-533      DataPointsFilter {filter, inPlaceFilter}    :485-497
//   OutlierFilter {compute}                       :544-553      ErrorMinimizer::ErrorElements               :742-755
//   ConvergenceError                              :139-143      Parametrizable::Parameters (string map)     Parametrizable.h:98-181
#pragma once
#include <cstddef>
#include <cstdint>
#include <limits>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace pmmirror {

// column-major dense matrix with the slice of Eigen's API the shims use
template <typename S>
class Mat {
 public:
  typedef std::ptrdiff_t Index;
  Mat() : r_(0), c_(0) {}
  Mat(Index r, Index c) : r_(r), c_(c), v_((size_t)r * (size_t)c) {}
  static Mat Zero(Index r, Index c) { Mat m(r, c); return m; }
  static Mat Constant(Index r, Index c, S val) { Mat m(r, c); for (auto& x : m.v_) x = val; return m; }
  static Mat Identity(Index r, Index c) { Mat m(r, c); for (Index i = 0; i < (r < c ? r : c); ++i) m(i, i) = S(1); return m; }
  Index rows() const { return r_; }
  Index cols() const { return c_; }
  S* data() { return v_.data(); }
  const S* data() const { return v_.data(); }
  S& operator()(Index i, Index j) { return v_[(size_t)i + (size_t)j * (size_t)r_]; }
  const S& operator()(Index i, Index j) const { return v_[(size_t)i + (size_t)j * (size_t)r_]; }
  void resize(Index r, Index c) { r_ = r; c_ = c; v_.assign((size_t)r * (size_t)c, S(0)); }
  void conservativeResize(Index r, Index c) {  // keeps the leading block, like Eigen
    Mat n(r, c);
    for (Index j = 0; j < (c < c_ ? c : c_); ++j)
      for (Index i = 0; i < (r < r_ ? r : r_); ++i) n(i, j) = (*this)(i, j);
    *this = n;
  }

 private:
  Index r_, c_;
  std::vector<S> v_;
};

struct PM {
  typedef float ScalarType;
  typedef Mat<float> Matrix;
  typedef Mat<int> IntMatrix;
  typedef Matrix TransformationParameters;
  typedef Matrix OutlierWeights;
  typedef std::map<std::string, std::string> Parameters;

  struct ConvergenceError : std::runtime_error { using std::runtime_error::runtime_error; };
  struct InvalidField : std::runtime_error { using std::runtime_error::runtime_error; };
  struct InvalidParameter : std::runtime_error { using std::runtime_error::runtime_error; };

  struct DataPoints {
    struct Label {
      std::string text;
      size_t span;
      Label(std::string t = "", size_t s = 0) : text(std::move(t)), span(s) {}
    };
    typedef std::vector<Label> Labels;
    Matrix features;       // (dim + 1) x N, homogeneous
    Labels featureLabels;
    Matrix descriptors;    // sum(span) x N
    Labels descriptorLabels;

    bool descriptorExists(const std::string& name) const {
      for (const auto& l : descriptorLabels) if (l.text == name) return true;
      return false;
    }
    // row offset of a descriptor block; throws InvalidField like getDescriptorViewByName (DataPoints.cpp)
    size_t descriptorStartingRow(const std::string& name, size_t* span = nullptr) const {
      size_t row = 0;
      for (const auto& l : descriptorLabels) {
        if (l.text == name) { if (span) *span = l.span; return row; }
        row += l.span;
      }
      throw InvalidField("Cannot find descriptor " + name);
    }
    // appends (or overwrites) a descriptor block, like DataPoints::addDescriptor
    void addDescriptor(const std::string& name, const Matrix& block) {
      if (descriptorExists(name)) {
        size_t span; const size_t row = descriptorStartingRow(name, &span);
        if ((size_t)block.rows() != span) throw InvalidField("descriptor " + name + " has another dimension");
        for (Matrix::Index j = 0; j < block.cols(); ++j)
          for (Matrix::Index i = 0; i < block.rows(); ++i) descriptors(row + i, j) = block(i, j);
        return;
      }
      const Matrix::Index old = descriptors.rows();
      Matrix n(old + block.rows(), block.cols());
      for (Matrix::Index j = 0; j < block.cols(); ++j) {
        for (Matrix::Index i = 0; i < old; ++i) n(i, j) = descriptors(i, j);
        for (Matrix::Index i = 0; i < block.rows(); ++i) n(old + i, j) = block(i, j);
      }
      descriptors = n;
      descriptorLabels.push_back(Label(name, (size_t)block.rows()));
    }
    unsigned getNbPoints() const { return (unsigned)features.cols(); }
  };

  struct Matches {
    typedef Matrix Dists;
    typedef IntMatrix Ids;
    static constexpr int InvalidId = -1;
    static constexpr float InvalidDist = std::numeric_limits<float>::infinity();
    Matches() {}
    Matches(const int knn, const int pointsCount) : dists(knn, pointsCount), ids(knn, pointsCount) {}
    Dists dists;
    Ids ids;
  };

  struct Parametrizable {
    Parameters parameters;
    Parametrizable() {}
    explicit Parametrizable(const Parameters& p) : parameters(p) {}
    virtual ~Parametrizable() {}
    std::string get(const std::string& name, const std::string& def) const {
      auto it = parameters.find(name);
      return it == parameters.end() ? def : it->second;
    }
  };

  struct Matcher : Parametrizable {
    unsigned long visitCounter = 0;
    using Parametrizable::Parametrizable;
    Matcher() {}
    virtual void init(const DataPoints& filteredReference) = 0;
    virtual Matches findClosests(const DataPoints& filteredReading) = 0;
  };

  struct DataPointsFilter : Parametrizable {
    using Parametrizable::Parametrizable;
    DataPointsFilter() {}
    virtual void init() {}
    virtual DataPoints filter(const DataPoints& input) = 0;
    virtual void inPlaceFilter(DataPoints& cloud) = 0;
  };

  struct OutlierFilter : Parametrizable {
    using Parametrizable::Parametrizable;
    OutlierFilter() {}
    virtual OutlierWeights compute(const DataPoints& filteredReading, const DataPoints& filteredReference,
                                   const Matches& input) = 0;
  };

  struct ErrorMinimizer : Parametrizable {
    struct ErrorElements {
      DataPoints reading, reference;
      OutlierWeights weights;
      Matches matches;
      int nbRejectedMatches = 0, nbRejectedPoints = 0;
      float pointUsedRatio = 0.f, weightedPointUsedRatio = 0.f;
    };
    using Parametrizable::Parametrizable;
    ErrorMinimizer() {}
    virtual TransformationParameters compute(const ErrorElements& matchedPoints) = 0;
  };
};

}  // namespace pmmirror
