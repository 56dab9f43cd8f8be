//! vec3.rs — `Vec3` with the reference's public surface (src/vec3.rs:3-47 and its operator impls).  `#[repr(C)]`:
//! a `&[Vec3]` is handed to the C ABI as `[B][3]` f32 without a copy.
use std::ops::{Add, Div, Mul, Sub};

#[repr(C)]
#[derive(Clone, Copy, Debug, PartialEq)]
pub struct Vec3 {
    pub x: f32,
    pub y: f32,
    pub z: f32,
}

impl Vec3 {
    pub fn new(x: f32, y: f32, z: f32) -> Self {
        Vec3 { x, y, z }
    }
    pub fn dot(&self, o: &Vec3) -> f32 {
        self.x * o.x + self.y * o.y + self.z * o.z
    }
    pub fn cross(&self, o: &Vec3) -> Vec3 {
        Vec3::new(self.y * o.z - self.z * o.y, self.z * o.x - self.x * o.z, self.x * o.y - self.y * o.x)
    }
    pub fn length(&self) -> f32 {
        self.dot(self).sqrt()
    }
    /// three divisions by the length, not a multiplication by its reciprocal (vec3.rs:31-34)
    pub fn normalize(&self) -> Vec3 {
        let l = self.length();
        Vec3::new(self.x / l, self.y / l, self.z / l)
    }
    pub fn scalar_multiply(&self, s: f32) -> Vec3 {
        Vec3::new(self.x * s, self.y * s, self.z * s)
    }
    pub fn scalar_divide(&self, s: f32) -> Vec3 {
        Vec3::new(self.x / s, self.y / s, self.z / s)
    }
    pub fn to_vec(&self) -> Vec<f32> {
        vec![self.x, self.y, self.z]
    }
    pub fn to_array(&self) -> [f32; 3] {
        [self.x, self.y, self.z]
    }
}

macro_rules! binop {
    ($tr:ident, $f:ident, $op:tt) => {
        impl $tr for Vec3 {
            type Output = Vec3;
            fn $f(self, r: Vec3) -> Vec3 {
                Vec3::new(self.x $op r.x, self.y $op r.y, self.z $op r.z)
            }
        }
        impl $tr<&Vec3> for Vec3 {
            type Output = Vec3;
            fn $f(self, r: &Vec3) -> Vec3 {
                Vec3::new(self.x $op r.x, self.y $op r.y, self.z $op r.z)
            }
        }
    };
}
binop!(Add, add, +);
binop!(Sub, sub, -);

impl Mul<f32> for Vec3 {
    type Output = Vec3;
    fn mul(self, s: f32) -> Vec3 {
        self.scalar_multiply(s)
    }
}
impl Div<f32> for Vec3 {
    type Output = Vec3;
    fn div(self, s: f32) -> Vec3 {
        self.scalar_divide(s)
    }
}
