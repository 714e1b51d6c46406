// pm_mirror.h — a LITERAL mirror of the parts of libpointmatcher's interface that the B200 shims
// (b200_pm_modules.hpp) touch, under the reference's OWN names (namespace PointMatcherSupport, template struct
// PointMatcher<T>), so that the shim header compiles unchanged against either this file or the reference's
// pointmatcher/PointMatcher.h (there is no #if between the two in b200_pm_modules.hpp). It exists only because this
// build image has no Eigen / Boost / yaml-cpp / libnabo, so the reference's own headers cannot be included here.
//
// Mirrored, with the reference's signatures, member names and behaviour (paths under libpointmatcher/pointmatcher/):
//   PointMatcherSupport::lexical_cast ("inf", "-inf", "nan")                         Parametrizable.h:57-91
//   PointMatcherSupport::Parametrizable {InvalidParameter, Comp<S>, ParameterDoc, ParametersDoc, Parameters,
//       ParametersUsed, className, parametersDoc, parameters, parametersUsed,
//       Parametrizable(className, paramsDoc, params) with the min/max checks, getParamValueString, get<S>}
//                                                                                   Parametrizable.h:102-181, .cpp:164-208
//   PointMatcherSupport::Registrar<Interface> {ClassDescriptor, GenericClassDescriptor<C> incl. the "was set but is
//       not used" check, GenericClassDescriptorNoParam<C>, reg, getDescriptor, create, getDescription, dump}, the
//       macros REG / DEF_REGISTRAR / ADD_TO_REGISTRAR / ADD_TO_REGISTRAR_NO_PARAM     Registrar.h:60-230
//       (createFromYAML needs yaml-cpp and is not mirrored: it is getNameParamsFromYAML + create(name, params))
//   PointMatcher<T>::{Matrix, IntMatrix, Vector, TransformationParameters, OutlierWeights}   PointMatcher.h:170-190
//   PointMatcher<T>::ConvergenceError :155-158      DataPoints {Label, Labels, InvalidField, features, descriptors,
//       descriptorExists, getDescriptorDimension, getDescriptorStartingRow, addDescriptor, getNbPoints}  :222-403
//   Matches :416-436   DataPointsFilter :485-497   Matcher :518-533   OutlierFilter :544-553
//   DegeneracyAwarenessMethod :576-583   LocalizabilityAnalysisResults :624-642
//   LocalizabilityParametersForErrorMinimization {A_, b_, deltas_, degeneracyAwarenessMethod, ...} :646-681
//   ErrorMinimizer {ErrorElements, compute(const ErrorElements&, LocalizabilityParametersForErrorMinimization&) = 0} :739-834
//   the four registrars the shims are added to (DEF_REGISTRAR) :478, 535, 566, 836
// Eigen matrices are stood in for by Mat<S>: column-major storage with the slice of Eigen's API the shims use
// (data(), rows(), cols(), operator()(i, j), resize, conservativeResize, Zero / Constant / Identity).
#pragma once
#include <cstddef>
#include <cstdint>
#include <iostream>
#include <limits>
#include <map>
#include <memory>
#include <set>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace PointMatcherSupport {

// ---- lexical_cast (Parametrizable.h:57-91; boost::lexical_cast semantics: the WHOLE string must parse) -----------
template <typename Target>
inline Target lexical_cast_plain(const std::string& arg) {
  std::istringstream is(arg);
  Target v;
  is >> std::noskipws >> v;
  if (is.fail() || is.peek() != std::char_traits<char>::eof()) throw std::runtime_error("bad lexical cast: " + arg);
  return v;
}
template <typename Target>
inline Target lexical_cast_scalar_to_string(const std::string& arg) {
  if (arg == "inf") return std::numeric_limits<Target>::infinity();
  if (arg == "-inf") return -std::numeric_limits<Target>::infinity();
  if (arg == "nan") return std::numeric_limits<Target>::quiet_NaN();
  return lexical_cast_plain<Target>(arg);
}
template <typename Target>
inline Target lexical_cast(const std::string& arg) { return lexical_cast_plain<Target>(arg); }
template <>
inline float lexical_cast<float>(const std::string& arg) { return lexical_cast_scalar_to_string<float>(arg); }
template <>
inline double lexical_cast<double>(const std::string& arg) { return lexical_cast_scalar_to_string<double>(arg); }
template <>
inline std::string lexical_cast<std::string>(const std::string& arg) { return arg; }
template <>
inline bool lexical_cast<bool>(const std::string& arg) {   // boost: exactly "0" or "1"
  if (arg == "1") return true;
  if (arg == "0") return false;
  throw std::runtime_error("bad lexical cast: " + arg);
}
template <typename S>
inline std::string toParam(const S& value) {
  std::ostringstream os;
  os << value;
  return os.str();
}

// ---- Parametrizable (Parametrizable.h:102-181, Parametrizable.cpp:43-208) -----------------------------------------
struct Parametrizable {
  struct InvalidParameter : std::runtime_error {
    InvalidParameter(const std::string& reason) : std::runtime_error(reason) {}
  };
  typedef bool (*LexicalComparison)(std::string a, std::string b);
  template <typename S>
  static bool Comp(std::string a, std::string b) { return lexical_cast<S>(a) < lexical_cast<S>(b); }
  static bool FalseLexicalComparison(std::string, std::string) { return false; }

  struct ParameterDoc {
    std::string name, doc, defaultValue, minValue, maxValue;
    LexicalComparison comp;
    ParameterDoc(const std::string& name, const std::string& doc, const std::string& defaultValue,
                 const std::string& minValue, const std::string& maxValue, LexicalComparison comp)
        : name(name), doc(doc), defaultValue(defaultValue), minValue(minValue), maxValue(maxValue), comp(comp) {}
    ParameterDoc(const std::string& name, const std::string& doc, const std::string& defaultValue)
        : name(name), doc(doc), defaultValue(defaultValue), minValue(""), maxValue(""), comp(FalseLexicalComparison) {}
  };
  typedef std::vector<ParameterDoc> ParametersDoc;
  typedef std::string Parameter;
  typedef std::map<std::string, Parameter> Parameters;
  typedef std::set<std::string> ParametersUsed;

  const std::string className;
  const ParametersDoc parametersDoc;
  Parameters parameters;
  ParametersUsed parametersUsed;

  Parametrizable() : className("unknown") {}
  Parametrizable(const std::string& className, const ParametersDoc paramsDoc, const Parameters& params)
      : className(className), parametersDoc(paramsDoc) {
    // fill current parameters from either values passed as argument, or default value (Parametrizable.cpp:171-189)
    for (auto it = parametersDoc.begin(); it != parametersDoc.end(); ++it) {
      const std::string& paramName(it->name);
      Parameters::const_iterator paramIt(params.find(paramName));
      if (paramIt != params.end()) {
        const std::string& val(paramIt->second);
        if (it->comp(val, it->minValue))
          throw InvalidParameter("Value " + val + " of parameter " + paramName + " in class " + className +
                                 " is smaller than minimum admissible value " + it->minValue);
        if (it->comp(it->maxValue, val))
          throw InvalidParameter("Value " + val + " of parameter " + paramName + " in class " + className +
                                 " is larger than maximum admissible value " + it->maxValue);
        parameters[paramName] = val;
      } else
        parameters[paramName] = it->defaultValue;
    }
  }
  virtual ~Parametrizable() {}

  std::string getParamValueString(const std::string& paramName) {
    Parameters::const_iterator paramIt(parameters.find(paramName));
    if (paramIt == parameters.end())
      throw InvalidParameter("Parameter " + paramName + " does not exist in class " + className);
    this->parametersUsed.insert(paramIt->first);
    return paramIt->second;
  }
  template <typename S>
  S get(const std::string& paramName) { return lexical_cast<S>(getParamValueString(paramName)); }
};

// ---- Registrar (Registrar.h:60-230) --------------------------------------------------------------------------------
struct InvalidElement : std::runtime_error {
  InvalidElement(const std::string& reason) : std::runtime_error(reason) {}
};

template <typename Interface>
struct Registrar {
 public:
  typedef Interface TargetType;
  struct ClassDescriptor {
    virtual ~ClassDescriptor() {}
    virtual std::shared_ptr<Interface> createInstance(const std::string& className,
                                                      const Parametrizable::Parameters& params) const = 0;
    virtual const std::string description() const = 0;
    virtual const Parametrizable::ParametersDoc availableParameters() const = 0;
  };
  template <typename C>
  struct GenericClassDescriptor : public ClassDescriptor {
    virtual std::shared_ptr<Interface> createInstance(const std::string& className,
                                                      const Parametrizable::Parameters& params) const {
      std::shared_ptr<C> instance = std::make_shared<C>(params);
      // check that all parameters were set (Registrar.h:103-111)
      for (auto it = params.begin(); it != params.end(); ++it) {
        if (instance->parametersUsed.find(it->first) == instance->parametersUsed.end())
          throw Parametrizable::InvalidParameter("Parameter " + it->first + " for module " + className +
                                                 " was set but is not used");
      }
      return instance;
    }
    virtual const std::string description() const { return C::description(); }
    virtual const Parametrizable::ParametersDoc availableParameters() const { return C::availableParameters(); }
  };
  template <typename C>
  struct GenericClassDescriptorNoParam : public ClassDescriptor {
    virtual std::shared_ptr<Interface> createInstance(const std::string& className,
                                                      const Parametrizable::Parameters& params) const {
      for (auto it = params.begin(); it != params.end(); ++it)
        throw Parametrizable::InvalidParameter("Parameter " + it->first + " was set but module " + className +
                                               " dos not use any parameter");
      return std::make_shared<C>();
    }
    virtual const std::string description() const { return C::description(); }
    virtual const Parametrizable::ParametersDoc availableParameters() const { return Parametrizable::ParametersDoc(); }
  };

 protected:
  typedef std::map<std::string, ClassDescriptor*> DescriptorMap;
  DescriptorMap classes;

 public:
  Registrar() {}
  Registrar(const Registrar&) = delete;
  Registrar& operator=(const Registrar&) = delete;
  ~Registrar() {
    for (auto it = classes.begin(); it != classes.end(); ++it) delete it->second;
  }
  void reg(const std::string& name, ClassDescriptor* descriptor) { classes[name] = descriptor; }
  const ClassDescriptor* getDescriptor(const std::string& name) const {
    auto it = classes.find(name);
    if (it == classes.end()) {
      std::cerr << "No element named " << name << " is registered. Known ones are:\n";
      dump(std::cerr);
      throw InvalidElement("Trying to instanciate unknown element " + name + " from registrar");
    }
    return it->second;
  }
  std::shared_ptr<Interface> create(const std::string& name,
                                    const Parametrizable::Parameters& params = Parametrizable::Parameters()) const {
    return getDescriptor(name)->createInstance(name, params);
  }
  const std::string getDescription(const std::string& name) const { return getDescriptor(name)->description(); }
  void dump(std::ostream& stream) const {
    for (auto it = classes.begin(); it != classes.end(); ++it) stream << "- " << it->first << "\n";
  }
  typename DescriptorMap::const_iterator begin() const { return classes.begin(); }
  typename DescriptorMap::const_iterator end() const { return classes.end(); }
};

#define REG(name) name##Registrar
#define DEF_REGISTRAR(name) PointMatcherSupport::Registrar<name> name##Registrar;
#define DEF_REGISTRAR_IFACE(name, ifaceName) PointMatcherSupport::Registrar<ifaceName> name##Registrar;
#define ADD_TO_REGISTRAR(name, elementName, element) { \
    typedef typename PointMatcherSupport::Registrar<name>::template GenericClassDescriptor<element> Desc; \
    name##Registrar.reg(#elementName, new Desc()); \
  }
#define ADD_TO_REGISTRAR_NO_PARAM(name, elementName, element) { \
    typedef typename PointMatcherSupport::Registrar<name>::template GenericClassDescriptorNoParam<element> Desc; \
    name##Registrar.reg(#elementName, new Desc()); \
  }

}  // namespace PointMatcherSupport

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

// ---- PointMatcher<T> (PointMatcher.h) --------------------------------------------------------------------------------
template <typename T>
struct PointMatcher {
  typedef T ScalarType;
  typedef Mat<T> Vector;                      // Eigen::Matrix<T, Dynamic, 1>: n x 1
  typedef Mat<T> Matrix;
  typedef Mat<int> IntMatrix;
  typedef Matrix TransformationParameters;
  typedef Matrix OutlierWeights;
  typedef PointMatcherSupport::Parametrizable Parametrizable;
  typedef Parametrizable::Parameters Parameters;
  typedef Parametrizable::ParameterDoc ParameterDoc;
  typedef Parametrizable::ParametersDoc ParametersDoc;
  typedef Parametrizable::InvalidParameter InvalidParameter;

  struct ConvergenceError : std::runtime_error {                                     // :155-158
    ConvergenceError(const std::string& reason) : std::runtime_error(reason) {}
  };

  struct DataPoints {                                                                // :222-403
    struct Label {
      std::string text;
      size_t span;
      Label(const std::string& text = "", const size_t span = 0) : text(text), span(span) {}
    };
    struct Labels : std::vector<Label> {
      Labels() {}
      Labels(const Label& label) { this->push_back(label); }
      bool contains(const std::string& text) const {
        for (const auto& l : *this) if (l.text == text) return true;
        return false;
      }
    };
    struct InvalidField : std::runtime_error {                                       // :280-283
      InvalidField(const std::string& reason) : std::runtime_error(reason) {}
    };

    Matrix features;       // (dim + 1) x N, homogeneous
    Labels featureLabels;
    Matrix descriptors;    // sum(span) x N
    Labels descriptorLabels;

    unsigned getNbPoints() const { return (unsigned)features.cols(); }
    bool descriptorExists(const std::string& name) const { return descriptorLabels.contains(name); }
    // getFieldDimension / getFieldStartingRow (DataPoints.cpp): 0 when the name is not found
    unsigned getDescriptorDimension(const std::string& name) const {
      for (const auto& l : descriptorLabels) if (l.text == name) return (unsigned)l.span;
      return 0;
    }
    unsigned getDescriptorStartingRow(const std::string& name) const {
      unsigned row = 0;
      for (const auto& l : descriptorLabels) {
        if (l.text == name) return row;
        row += (unsigned)l.span;
      }
      return 0;
    }
    // addDescriptor -> addField (DataPoints.cpp): overwrite a block of the same span, else append
    void addDescriptor(const std::string& name, const Matrix& newDescriptor) {
      if (descriptorExists(name)) {
        const unsigned span = getDescriptorDimension(name), row = getDescriptorStartingRow(name);
        if ((unsigned)newDescriptor.rows() != span)
          throw InvalidField("The existing field " + name + " has dimension different than the new one");
        for (typename Matrix::Index j = 0; j < newDescriptor.cols(); ++j)
          for (typename Matrix::Index i = 0; i < newDescriptor.rows(); ++i) descriptors(row + i, j) = newDescriptor(i, j);
        return;
      }
      if (descriptors.cols() != 0 && newDescriptor.cols() != descriptors.cols())
        throw InvalidField("Wrong field dimension");
      const typename Matrix::Index old = descriptors.rows();
      Matrix n(old + newDescriptor.rows(), newDescriptor.cols());
      for (typename Matrix::Index j = 0; j < newDescriptor.cols(); ++j) {
        for (typename Matrix::Index i = 0; i < old; ++i) n(i, j) = descriptors(i, j);
        for (typename Matrix::Index i = 0; i < newDescriptor.rows(); ++i) n(old + i, j) = newDescriptor(i, j);
      }
      descriptors = n;
      descriptorLabels.push_back(Label(name, (size_t)newDescriptor.rows()));
    }
  };

  struct Matches {                                                                   // :416-436
    typedef Matrix Dists;
    typedef IntMatrix Ids;
    static constexpr int InvalidId = -1;
    static constexpr T InvalidDist = std::numeric_limits<T>::infinity();
    Matches() {}
    Matches(const Dists& dists, const Ids ids) : dists(dists), ids(ids) {}
    Matches(const int knn, const int pointsCount) : dists(knn, pointsCount), ids(knn, pointsCount) {}
    Dists dists;
    Ids ids;
  };

  struct DataPointsFilter : public Parametrizable {                                  // :485-497
    DataPointsFilter() {}
    DataPointsFilter(const std::string& className, const ParametersDoc paramsDoc, const Parameters& params)
        : Parametrizable(className, paramsDoc, params) {}
    virtual ~DataPointsFilter() {}
    virtual void init() {}
    virtual DataPoints filter(const DataPoints& input) = 0;
    virtual void inPlaceFilter(DataPoints& cloud) = 0;
  };
  DEF_REGISTRAR(DataPointsFilter)                                                    // :509

  struct Matcher : public Parametrizable {                                           // :518-533
    unsigned long visitCounter;
    Matcher() : visitCounter(0) {}
    Matcher(const std::string& className, const ParametersDoc paramsDoc, const Parameters& params)
        : Parametrizable(className, paramsDoc, params), visitCounter(0) {}
    virtual ~Matcher() {}
    void resetVisitCount() { visitCounter = 0; }
    unsigned long getVisitCount() const { return visitCounter; }
    virtual void init(const DataPoints& filteredReference) = 0;
    virtual Matches findClosests(const DataPoints& filteredReading) = 0;
  };
  DEF_REGISTRAR(Matcher)                                                             // :535

  struct OutlierFilter : public Parametrizable {                                     // :544-553
    OutlierFilter() {}
    OutlierFilter(const std::string& className, const ParametersDoc paramsDoc, const Parameters& params)
        : Parametrizable(className, paramsDoc, params) {}
    virtual ~OutlierFilter() {}
    virtual OutlierWeights compute(const DataPoints& filteredReading, const DataPoints& filteredReference,
                                   const Matches& input) = 0;
  };
  DEF_REGISTRAR(OutlierFilter)                                                       // :566

  enum class DegeneracyAwarenessMethod {                                             // :576-583
    kNone = 0, kSolutionRemapping = 1, kOptimizedEqualityConstraints = 2, kEqualityConstraints = 3,
    kInequalityConstraints = 4
  };
  enum class LocalizabilityCategory { kNonLocalizable = 0, kLocalizable = 1 };        // :597-601
  struct LocalizabilityConstraints {                                                 // :617-621
    Vector translationConstraintValues_ = Vector::Zero(3, 1);
    Vector rotationConstraintValues_ = Vector::Zero(3, 1);
  };
  struct LocalizabilityAnalysisResults {                                             // :624-642
    Matrix translationEigenvectors_ = Matrix::Zero(3, 3);
    Matrix rotationEigenvectors_ = Matrix::Zero(3, 3);
    Vector localizabilityXyz_ = Vector::Constant(3, 1, static_cast<T>(LocalizabilityCategory::kLocalizable));
    Vector localizabilityRpy_ = Vector::Constant(3, 1, static_cast<T>(LocalizabilityCategory::kLocalizable));
    LocalizabilityConstraints localizabilityConstraints_;
    Matrix solutionRemappingProjectionMatrix_ = Matrix::Identity(6, 6);
  };
  struct LocalizabilityParametersForErrorMinimization {                              // :646-681
    Matrix A_ = Matrix::Zero(6, 6);     // Eigen::Matrix<T, 6, 6>: optimisation Hessian formed by
    Matrix b_ = Matrix::Zero(6, 1);     // ICP::calculateOptimizationHessian (ICP.cpp:1513-1566) before compute()
    Matrix deltas_;
    DegeneracyAwarenessMethod degeneracyAwarenessMethod{DegeneracyAwarenessMethod::kNone};
    LocalizabilityAnalysisResults localizabilityAnalysisResults_;
    TransformationParameters T_convert_from_base_to_lidar_mat;
    std::vector<double> residuals_, regNorms_, lambdas_;
    int iterationNumber_{0};
  };

  struct ErrorMinimizer : public Parametrizable {                                    // :739-834
    struct ErrorElements {
      DataPoints reading, reference;
      OutlierWeights weights;
      Matches matches;
      int nbRejectedMatches, nbRejectedPoints;
      T pointUsedRatio, weightedPointUsedRatio;
      ErrorElements() : nbRejectedMatches(-1), nbRejectedPoints(-1), pointUsedRatio(-1.), weightedPointUsedRatio(-1.) {}
    };
    ErrorMinimizer() {}
    ErrorMinimizer(const std::string& className, const ParametersDoc paramsDoc, const Parameters& params)
        : Parametrizable(className, paramsDoc, params) {}
    virtual ~ErrorMinimizer() {}
    //! Find the transformation that minimizes the error given matched pair of points (:790, pure virtual)
    virtual TransformationParameters compute(const ErrorElements& matchedPoints,
                                             LocalizabilityParametersForErrorMinimization& localizabilityParametersForErrorMinimization) = 0;
  };
  DEF_REGISTRAR(ErrorMinimizer)                                                      // :836

  // the reference's constructor (Registry.cpp:61-128) registers every stock module here; the B200 lines are the ones
  // INTEGRATION.md adds (host/selftest.cpp adds them to an instance of this mirror)
  PointMatcher() {}
};
