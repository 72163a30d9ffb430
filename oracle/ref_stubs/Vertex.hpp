// TEST INFRASTRUCTURE ONLY — stand-in for UniEngine's Vertex.hpp (another repository, absent here).  Member names and
// order are the ones the reference's own code uses (RayDataDefinations.hpp:132-161: m_position / m_positionPadding /
// m_normal / m_normalPadding / m_tangent / m_tangentPadding / m_color / m_texCoord / m_texCoordPadding; RT.cu:1163-1173),
// 80 bytes.  SkinnedVertex is read by ref_prebake_driver.cpp only; StrandPoint is only stored in host containers, except that
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
// Flat, as CopySkinnedVerticesKernel reads it (RT.cu:1194-1247: m_position, m_bondId[k], m_weight[k], m_bondId2[k],
// m_weight2[k], the four paddings); member order = rtf_skinned_vertex (include/rtf/rtf.h), 144 bytes.
struct SkinnedVertex {
  glm::vec3 m_position;
  float m_positionPadding = 0.0f;
  glm::vec3 m_normal;
  float m_normalPadding = 0.0f;
  glm::vec3 m_tangent;
  float m_tangentPadding = 0.0f;
  glm::vec4 m_color;
  glm::vec2 m_texCoord;
  glm::vec2 m_texCoordPadding;
  int m_bondId[4];
  glm::vec4 m_weight;
  int m_bondId2[4];
  glm::vec4 m_weight2;
};
struct StrandPoint {
  glm::vec3 m_position;
  float m_thickness = 0.0f;
  glm::vec4 m_color;
  glm::vec3 m_normal;
  float m_texCoord = 0.0f;
};
}  // namespace UniEngine
