"""CPU: the mesh self-oracle (oracle/icx_oracle.cpp icxo_mesh_pass) against answers known in closed form.  The reference has
no mesh collider, so this is all the pinning the extension can get: normals, depths and contact points of simple poses, the
ascending-triangle order, and the response (static resolution moves the body out by exactly the depth)."""
import numpy as np

from intercession_b200 import abi, scenes
import refs


def one_body(kind, origin, scale=1.0, velocity=(0, -1, 0)):
    s = scenes._empty(1)
    s["ids"] = np.array([abi.compose_entity(0, 1, abi.NULL_CHAINLINK)], np.uint32)
    s.T["origin"][0] = origin
    s.T["scale"][0] = scale
    s.P["mass"][0] = 2.0
    s.P["velocity"][0] = velocity
    c = s.C["colliders"]
    c["colliderType"][0, 0] = kind
    c["isActive"][0, 0] = 1
    c["isActive"][0, 1] = 0
    c["collisionType"][0, 0] = abi.COLLISION_RIGID
    return s


FLOOR_V = np.array([[-5, 0, -5], [5, 0, -5], [5, 0, 5], [-5, 0, 5]], np.float32)
FLOOR_T = np.array([[0, 2, 1], [0, 3, 2]], np.uint32)                       # both wound so that the geometric normal is +y


def test_sphere_on_a_flat_triangle():
    s = one_body(abi.COLLIDER_SPHERE, (1.0, 0.3, -2.0))                     # radius 0.5 (unit diameter x scale 1), centre 0.3 above: depth 0.2
    out, hits = refs.oracle_mesh_pass(s, FLOOR_V, FLOOR_T)
    assert len(hits) == 1 and hits["triangle"][0] == 0                      # the point lies in triangle 0 only; after the push-out triangle 1 is touched, not penetrated
    assert np.allclose(hits["normal"][0], (0, 1, 0), atol=2e-5) and np.isclose(hits["depth"][0], 0.2) and np.allclose(hits["point"][0], (1.0, 0.0, -2.0), atol=2e-5)
    assert np.isclose(out.T["origin"][0, 1], 0.5)                            # static resolution: out by exactly the depth (the mesh does not move)
    assert out.P["velocity"][0, 1] > 0                                       # restitution 0.7 * 0.7: it bounces


def test_sphere_against_an_edge_and_a_miss():
    s = one_body(abi.COLLIDER_SPHERE, (5.3, 0.0, 0.0))                      # beside the floor's x = 5 edge: closest point on the edge
    _, hits = refs.oracle_mesh_pass(s, FLOOR_V, FLOOR_T)
    assert len(hits) >= 1 and np.allclose(hits["normal"][0], (1, 0, 0), atol=2e-5) and np.isclose(hits["depth"][0], 0.2, atol=1e-6)
    assert np.allclose(hits["point"][0], (5.0, 0.0, 0.0), atol=2e-5)
    _, none = refs.oracle_mesh_pass(one_body(abi.COLLIDER_SPHERE, (0, 0.6, 0)), FLOOR_V, FLOOR_T)
    assert len(none) == 0


def test_box_resting_and_tilted():
    s = one_body(abi.COLLIDER_BOX, (0.5, 0.4, -1.0))                         # unit cube, centre 0.4 above the floor: depth 0.1 along +y
    out, hits = refs.oracle_mesh_pass(s, FLOOR_V, FLOOR_T)
    assert len(hits) >= 1 and np.allclose(hits["normal"][0], (0, 1, 0), atol=2e-5) and np.isclose(hits["depth"][0], 0.1, atol=1e-6)
    assert np.allclose(hits["point"][0], (0.5, 0.0, -1.0), atol=2e-5)                   # the triangle's closest point to the box centre
    assert np.isclose(out.T["origin"][0, 1], 0.5, atol=1e-6)
    # below the sheet the same box is pushed out downwards (two-sided surface)
    _, under = refs.oracle_mesh_pass(one_body(abi.COLLIDER_BOX, (0.5, -0.4, -1.0)), FLOOR_V, FLOOR_T)
    assert np.allclose(under["normal"][0], (0, -1, 0), atol=2e-5)
    # rotated 45 degrees about z: the lowest edge is sqrt(0.5) below the centre
    t = one_body(abi.COLLIDER_BOX, (0.0, 0.6, 0.0))
    t.T["orientation"][0] = (np.cos(np.pi / 8), 0, 0, np.sin(np.pi / 8))
    _, tilted = refs.oracle_mesh_pass(t, FLOOR_V, FLOOR_T)
    assert np.allclose(tilted["normal"][0], (0, 1, 0), atol=1e-6) and np.isclose(tilted["depth"][0], np.sqrt(0.5) - 0.6, atol=1e-5)


def test_triangles_are_met_in_ascending_order_and_later_ones_see_the_moved_body():
    # two overlapping sheets: y = 0 (triangles 0, 1) and y = 0.1 (triangles 2, 3); a sphere at 0.3 penetrates both
    v = np.concatenate([FLOOR_V, FLOOR_V + np.float32([0, 0.1, 0])])
    t = np.concatenate([FLOOR_T, FLOOR_T + 4]).astype(np.uint32)
    s = one_body(abi.COLLIDER_SPHERE, (1.0, 0.3, -2.0))
    out, hits = refs.oracle_mesh_pass(s, v, t)
    assert hits["triangle"].tolist() == [0, 2]
    assert np.isclose(hits["depth"][0], 0.2) and np.isclose(hits["depth"][1], 0.1, atol=1e-6)      # 0.5 - (0.5 - 0.1): pushed to 0.5 by the first
    assert np.isclose(out.T["origin"][0, 1], 0.6, atol=1e-6)


def test_ray_spring_on_a_flat_triangle():
    """The biped's legs (a ray collider of collision type spring, client_focal_entity.h:30-91) over mesh ground."""
    h = np.float32(np.sqrt(0.5))
    s = one_body(abi.COLLIDER_BOX, (1.0, 2.0, -2.0))
    c = s.C["colliders"]
    c["isActive"][0, 1] = 1
    c["colliderType"][0, 1] = abi.COLLIDER_RAY
    c["collisionType"][0, 1] = abi.COLLISION_SPRING
    c["localTransform"]["orientation"][0, 1] = (h, h, 0, 0)                   # +z of the ray -> -y of the body
    c["localTransform"]["scale"][0, 1] = (1, 1, 3.0)                          # 3 long: the end is 1 below the floor
    c["stiffness"][0, 1], c["damping"][0, 1], c["restLength"][0, 1] = 2500, 400, 0.1
    c["inheritOrientation"][0, 1], c["influenceOrientation"][0, 1] = 0, 0
    c["minParametricValue"][0, 1] = 1.0
    out, hits = refs.oracle_mesh_pass(s, FLOOR_V, FLOOR_T)
    ray = hits[hits["slot"] == 1]
    assert len(ray) == 1                                                      # the second floor triangle lies no nearer: minParametricValue keeps the first
    assert np.allclose(ray["normal"][0], (0, 1, 0), atol=2e-5) and np.isclose(ray["depth"][0], 1.0, atol=1e-5)
    assert np.allclose(ray["point"][0], (1.0, -0.5, -2.0), atol=2e-5)          # spring_rigid_response moves the contact half the depth in (:926)
    assert np.isclose(out.C["colliders"]["minParametricValue"][0, 1], 2.0 / 3.0, atol=1e-6)
    assert out.P["acceleration"][0, 1] > 0                                    # the spring pushes the body up
    # a ray that ends above the floor meets nothing; a hit behind an earlier one (minParametricValue) is not taken
    c["localTransform"]["scale"][0, 1] = (1, 1, 1.5)
    assert not (refs.oracle_mesh_pass(s, FLOOR_V, FLOOR_T)[1]["slot"] == 1).any()
    c["localTransform"]["scale"][0, 1] = (1, 1, 3.0)
    c["minParametricValue"][0, 1] = 0.5
    assert not (refs.oracle_mesh_pass(s, FLOOR_V, FLOOR_T)[1]["slot"] == 1).any()
