// Stand-in for the mrpt::opengl scene-graph classes libopenbeam's getVisualization() methods build.  There is no
// graphics stack in this environment (and the renderers are out of scope): the objects accept the calls and do nothing.
#pragma once
#include <memory>
#include <string>
#include <mrpt/math/TPoint3D.h>
namespace mrpt::img {
struct TColor {
  unsigned char R = 0, G = 0, B = 0, A = 255;
  static TColor blue() { return TColor{0, 0, 255, 255}; }
};
}  // namespace mrpt::img
namespace mrpt::poses {
struct CPose3D {
  CPose3D() = default;
  CPose3D(double, double, double, double = 0, double = 0, double = 0) {}
};
}  // namespace mrpt::poses
namespace mrpt::opengl {
struct CRenderizable {
  using Ptr = std::shared_ptr<CRenderizable>;
  virtual ~CRenderizable() = default;
  template <class... A> void setColor(A&&...) {}
  template <class... A> void setColor_u8(A&&...) {}
  template <class... A> void setLocation(A&&...) {}
  template <class... A> void setPose(A&&...) {}
  template <class... A> void setName(A&&...) {}
};
template <class D> struct Creatable : CRenderizable {
  using Ptr = std::shared_ptr<D>;
  static Ptr Create() { return std::make_shared<D>(); }
};
struct CSetOfObjects : Creatable<CSetOfObjects> {
  template <class P> void insert(const P&) {}
  void clear() {}
};
struct CArrow : Creatable<CArrow> {
  template <class... A> void setArrowEnds(A&&...) {}
  void setHeadRatio(double) {}
  void setSmallRadius(double) {}
  void setLargeRadius(double) {}
};
struct CSphere : Creatable<CSphere> {
  void setRadius(double) {}
  void setNumberDivsLatitude(int) {}
  void setNumberDivsLongitude(int) {}
};
struct CCylinder : Creatable<CCylinder> {
  template <class... A> void setRadii(A&&...) {}
  void setHeight(double) {}
};
struct CText : Creatable<CText> {
  void setString(const std::string&) {}
};
struct CSetOfLines : Creatable<CSetOfLines> {
  template <class... A> void appendLine(A&&...) {}
  void setLineWidth(double) {}
};
}  // namespace mrpt::opengl
