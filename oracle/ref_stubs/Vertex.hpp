// TEST INFRASTRUCTURE ONLY — stand-in for UniEngine's Vertex.hpp (another repository, absent here).  Member names and
// order are the ones the reference's own code uses (RayDataDefinations.hpp:132-161: m_position / m_positionPadding /
// m_normal / m_normalPadding / m_tangent / m_tangentPadding / m_color / m_texCoord / m_texCoordPadding; RT.cu:1163-1173),
// 80 bytes.  SkinnedVertex / StrandPoint are only stored in host containers the host drivers never touch, except that
// CurveSplineDefinations.hpp reads m_position / m_thickness / m_color / m_texCoord of a StrandPoint.
#pragma once
#include <glm/glm.hpp>
namespace UniEngine {
struct Vertex {
  glm::vec3 m_position;
  float m_positionPadding = 0.0f;
  glm::vec3 m_normal;
  float m_normalPadding = 0.0f;
  glm::vec3 m_tangent;
  float m_tangentPadding = 0.0f;
  glm::vec4 m_color;
  glm::vec2 m_texCoord;
  glm::vec2 m_texCoordPadding;
};
struct SkinnedVertex {
  Vertex m_v;
  int m_bondId[4], m_bondId2[4];
  glm::vec4 m_weight, m_weight2;
};
struct StrandPoint {
  glm::vec3 m_position;
  float m_thickness = 0.0f;
  glm::vec4 m_color;
  glm::vec3 m_normal;
  float m_texCoord = 0.0f;
};
}  // namespace UniEngine
