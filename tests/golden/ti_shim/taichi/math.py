"""`taichi.math` of the shim: vecN / matN constructors and the few functions the hot path uses."""
from ._scalar import cos, exp, floor, log, sin, sqrt  # noqa: F401
from .matrix import Matrix, make_matrix_type, make_vector_type

vec2 = make_vector_type(2)
vec3 = make_vector_type(3)
vec4 = make_vector_type(4)
mat2 = make_matrix_type(2, 2)
mat3 = make_matrix_type(3, 3)
mat4 = make_matrix_type(4, 4)
pi = 3.141592653589793


def dot(a, b):
    return a.dot(b)


def normalize(v):
    return v.normalized()


def length(v):
    return v.norm()


def cross(a, b):
    return Matrix([a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]])
