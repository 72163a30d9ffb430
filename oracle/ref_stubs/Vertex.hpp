// TEST INFRASTRUCTURE ONLY — stand-in for UniEngine's Vertex.hpp (another repository, absent here): RayTracer.hpp only
// stores these types in host containers the host drivers never touch.  Field lists follow the 80 / 144 / 48 byte
// layouts SURVEY.md §8 records; nothing depends on them.
#pragma once
#include <glm/glm.hpp>
namespace UniEngine {
struct Vertex {
  glm::vec3 m_position, m_normal, m_tangent;
  glm::vec4 m_color;
  glm::vec2 m_texCoord;
  glm::vec4 m_positionPadding;
};
struct SkinnedVertex {
  Vertex m_v;
  int m_bondId[4], m_bondId2[4];
  glm::vec4 m_weight, m_weight2;
};
struct StrandPoint {
  glm::vec3 m_position;
  float m_thickness;
  glm::vec4 m_color;
  glm::vec3 m_normal;
  float m_texCoord;
};
}  // namespace UniEngine
