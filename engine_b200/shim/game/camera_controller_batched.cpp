// See camera_controller_batched.hpp.  Arithmetic follows the reference expression by expression (same glm calls,
// same operand order) so that a trajectory is bit-identical to the original controller's.
#include "camera_controller_batched.hpp"

#define GLM_ENABLE_EXPERIMENTAL
#include <array>
#include <cmath>
#include <limits>
#include <typeinfo>

#include <glm/gtc/constants.hpp>
#include <glm/gtc/quaternion.hpp>
#include <glm/gtx/rotate_vector.hpp>

#include "application.h"
#include "input_manager.h"
#include "scene.h"

namespace {
using Cfg = CameraControllerComponent;
constexpr std::size_t kWallRays = Cfg::kNumHorizontalRays;

// reference :83-90 — probe i of the wall fan: feet-to-eyes spread along -z, direction = horizontal velocity
engine::Ray WallProbe(const glm::vec3& eye, const glm::vec3& vel, std::size_t i)
{
    const float lerp_value = static_cast<float>(i) / (static_cast<float>(kWallRays) - 1);
    engine::Ray ray{};
    ray.origin = eye;
    ray.origin.z -= (Cfg::kPlayerHeight - Cfg::kMaxStairHeight) * lerp_value;
    ray.direction.x = vel.x;
    ray.direction.y = vel.y;  // GetRaycast normalises
    ray.direction.z = 0.0f;
    return ray;
}

// reference :120-123 (falling) / :146-148 (rising)
engine::Ray VerticalProbe(const glm::vec3& eye, bool falling)
{
    engine::Ray ray{};
    ray.origin = eye;
    if (falling) {
        ray.origin.z += Cfg::kMaxStairHeight - Cfg::kPlayerHeight;
        ray.direction = {0.0f, 0.0f, -1.0f};
    }
    else {
        ray.direction = {0.0f, 0.0f, 1.0f};
    }
    return ray;
}
}  // namespace

BatchedCameraControllerSystem::BatchedCameraControllerSystem(engine::Scene* scene)
    : System(scene, {typeid(engine::TransformComponent).hash_code(), typeid(CameraControllerComponent).hash_code()})
{
}

bool BatchedCameraControllerSystem::Bind()
{
    if (transform_ && state_) return true;
    for (engine::Entity entity : m_entities) {  // the first (only) controlled entity, as the reference does (:29-33)
        transform_ = m_scene->GetComponent<engine::TransformComponent>(entity);
        state_ = m_scene->GetComponent<CameraControllerComponent>(entity);
        break;
    }
    collisions_ = m_scene->GetSystem<engine::CollisionSystem>();
    return transform_ && state_;
}

void BatchedCameraControllerSystem::onUpdate(float ts)
{
    if (!Bind()) return;
    engine::TransformComponent& t = *transform_;
    CameraControllerComponent& c = *state_;
    engine::InputManager* input = m_scene->app()->getInputManager();
    const float dt = ts;
    ++frames;

    // ---- input -> look angles and velocity (reference :40-70) -------------------------------------------------
    float strafe = input->GetAxis("movex") * Cfg::kSpeedStrafe;
    float advance = input->GetAxis("movey") * Cfg::kSpeedForwardBack;
    if (input->GetButton("sprint")) advance *= Cfg::kSprintMultiplier;

    const float d_pitch = input->GetAxis("looky") * -1.0f * Cfg::kCameraSensitivity;
    c.pitch += d_pitch;
    if (c.pitch <= Cfg::kMinPitch || c.pitch >= Cfg::kMaxPitch) c.pitch -= d_pitch;
    c.yaw += input->GetAxis("lookx") * -1.0f * Cfg::kCameraSensitivity;

    const glm::vec3 side_step = glm::rotateZ(glm::vec3{strafe, 0.0f, 0.0f}, c.yaw);
    const glm::vec3 fwd_step = glm::rotateZ(glm::vec3{0.0f, advance, 0.0f}, c.yaw);
    const glm::vec3 planar = (side_step + fwd_step);
    c.vel.x = planar.x;
    c.vel.y = planar.y;
    c.vel.z += c.grav_accel * dt;
    if (input->GetButtonPress("jump") && (c.grounded || c.noclip)) c.vel.z += Cfg::kJumpVelocity;

    // ---- this frame's probes, one batch -------------------------------------------------------------------------
    const bool want_walls = (c.vel.x != 0.0f || c.vel.y != 0.0f) && !c.noclip;  // :75
    const bool falling = c.vel.z < 0.0f, rising = c.vel.z > 0.0f;              // :119, :143
    const bool want_vertical = !c.noclip && (falling || rising);
    std::array<engine::Ray, kWallRays + 1> rays{};
    std::array<engine::Raycast, kWallRays + 1> hits{};
    std::size_t n = 0;
    if (want_walls)
        for (std::size_t i = 0; i < kWallRays; ++i) rays[n++] = WallProbe(t.position, c.vel, i);
    const std::size_t vertical_slot = n;
    if (want_vertical) rays[n++] = VerticalProbe(t.position, falling);  // speculative: from the unpushed position
    if (n > 0) {
        collisions_->GetRaycasts(rays.data(), n, hits.data());
        ++raycast_calls;
        rays_cast += n;
    }

    // ---- walls (reference :77-114) ------------------------------------------------------------------------------
    if (want_walls) {
        const engine::Raycast* nearest = nullptr;
        float nearest_distance = std::numeric_limits<float>::infinity();
        for (std::size_t i = 0; i < kWallRays; ++i) {
            if (hits[i].hit && hits[i].distance < nearest_distance) {
                nearest_distance = hits[i].distance;
                nearest = &hits[i];
            }
        }
        if (nearest != nullptr) {
            const glm::vec2 into_wall = glm::normalize(glm::vec2{nearest->normal.x, nearest->normal.y}) * -1.0f;
            const glm::vec2 vel_xy = glm::vec2{c.vel.x, c.vel.y};
            const glm::vec2 vel_into_wall = into_wall * glm::dot(into_wall, vel_xy);
            const glm::vec2 step_into_wall = vel_into_wall * dt;
            if (glm::length(step_into_wall) > nearest->distance - Cfg::kPlayerCollisionRadius) {
                const glm::vec2 back_off = glm::normalize(vel_xy) * fmaxf(Cfg::kPlayerCollisionRadius, nearest->distance);
                const glm::vec3 before = t.position;
                t.position.x = nearest->location.x - back_off.x;
                t.position.y = nearest->location.y - back_off.y;
                c.vel.x -= vel_into_wall.x;
                c.vel.y -= vel_into_wall.y;
                if (want_vertical && (t.position.x != before.x || t.position.y != before.y)) {
                    // the speculative vertical probe started from the wrong column: cast it again, alone
                    rays[vertical_slot] = VerticalProbe(t.position, falling);
                    collisions_->GetRaycasts(&rays[vertical_slot], 1, &hits[vertical_slot]);
                    ++raycast_calls;
                    ++rays_cast;
                    ++recasts;
                }
            }
        }
    }

    // ---- floor / ceiling (reference :118-161) -------------------------------------------------------------------
    if (want_vertical) {
        const engine::Raycast& v = hits[vertical_slot];
        if (falling) {
            c.grounded = false;
            if (v.hit) {
                const float drop = fabsf(c.vel.z * dt);
                if (drop > v.distance - Cfg::kMaxStairHeight) {  // would end up inside the step: stand on it
                    t.position.z = v.location.z + Cfg::kPlayerHeight;
                    c.vel.z = 0.0f;
                    c.grounded = true;
                }
            }
        }
        else {
            c.grounded = false;
            if (v.hit) {
                const float rise = fabsf(c.vel.z * dt);
                if (rise > v.distance - Cfg::kPlayerCollisionRadius) {
                    t.position.z = v.location.z - Cfg::kPlayerCollisionRadius;
                    c.vel.z = 0.0f;
                }
            }
        }
    }

    t.position += c.vel * dt;  // :163

    // ---- orientation (reference :167-184): yaw about z times pitch about x --------------------------------------
    const float half_pitch = c.pitch / 2.0f, half_yaw = c.yaw / 2.0f;
    glm::quat about_x{};
    about_x.x = glm::sin(half_pitch);
    about_x.y = 0.0f;
    about_x.z = 0.0f;
    about_x.w = glm::cos(half_pitch);
    glm::quat about_z{};
    about_z.x = 0.0f;
    about_z.y = 0.0f;
    about_z.z = glm::sin(half_yaw);
    about_z.w = glm::cos(half_yaw);
    t.rotation = about_z * about_x;

    // ---- fell out of the world (reference :188-193, without the reset key) --------------------------------------
    if (glm::length(t.position) > Cfg::kMaxDistanceFromOrigin) {
        t.position = {0.000001f, 0.000001f, 10.0f};
        c.vel = {0.0f, 0.0f, 0.0f};
        c.pitch = glm::half_pi<float>();
        c.yaw = 0.0f;
    }
}
