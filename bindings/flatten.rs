// UNBUILT here: this image has no rustc/cargo (one statement of that for the whole directory: see
// INTEGRATION.md §0).  Kept in sync with include/rrt.h by tests/test_abi.py.
//
// src/rrt_flatten.rs — walks the reference's `HittableObject` tree once and emits the flat arrays of
// `rrt_scene_desc` (include/rrt.h): one `RrtHittable` per `Arc<HittableObject>` in construction order,
// shared nodes emitted once, materials / textures / images / Perlin tables de-duplicated by pointer.
// It replaces `gpu::build_in_one_weekend_scene`'s hand-written POD conversion (src/gpu/mod.rs:124-301),
// which only knew spheres and three materials.
//
// What the reference's structs must expose for this visitor (they are private fields today; the
// visitor lives in the same crate, so `pub(crate)` is enough):
//   Sphere { center: Ray, radius, mat }                     the_rest_of_your_life/sphere.rs:10-15
//   Quad { q, u, v, mat }                                   quad.rs:10-20
//   Translate { object, offset }, RotateY { object, .. }    hittable.rs:73-105
//   ConstantMedium { boundary, neg_inv_density, phase_function }   constant_medium.rs:10-14
//   Lambertian/DiffuseLight/Isotropic { tex }, Metal { albedo, fuzz }, Dielectric { refraction_index }
//   SolidColor { albedo }, CheckerTexture { inv_scale, even, odd }, ImageTexture { image },
//   NoiseTexture { noise, scale }, Perlin { randvec, perm_x, perm_y, perm_z }, RtwImage { data, width, height }
// and two values they discard today and must keep:
//   BvhNode.source_objects: Vec<HittableRef>   the list BvhNode::new was given (bvh.rs:16-19) — the library
//                                              builds its own tree and replays the reference's split order
//                                              only to rank exact-t ties;
//   RotateY.angle_degrees: f64                 hittable.rs:107-112 keeps only sin/cos.
use std::collections::HashMap;
use std::sync::Arc;

use crate::books::the_rest_of_your_life::hittable::{HittableObject, HittableRef};
use crate::books::the_rest_of_your_life::material::{MaterialObject, MaterialRef};
use crate::books::the_rest_of_your_life::perlin::Perlin;
use crate::books::the_rest_of_your_life::rtw_image::RtwImage;
use crate::books::the_rest_of_your_life::texture::{TextureObject, TextureRef};
use crate::rrt_ffi::*;

// include/rrt.h enums
const H_SPHERE: i32 = 0;
const H_QUAD: i32 = 1;
const H_LIST: i32 = 2;
const H_BVH: i32 = 3;
const H_TRANSLATE: i32 = 4;
const H_ROTATE_Y: i32 = 5;
const H_CONSTANT_MEDIUM: i32 = 6;
const M_LAMBERTIAN: i32 = 0;
const M_METAL: i32 = 1;
const M_DIELECTRIC: i32 = 2;
const M_DIFFUSE_LIGHT: i32 = 3;
const M_ISOTROPIC: i32 = 4;
const M_EMPTY: i32 = 5;
const T_SOLID: i32 = 0;
const T_CHECKER: i32 = 1;
const T_IMAGE: i32 = 2;
const T_NOISE: i32 = 3;

/// Owns every array an `RrtSceneDesc` points into.  The description borrows from it, so it must
/// outlive the `rrt_scene_create` / `rrt_render_frame` call (the library copies everything during
/// that call and keeps no host pointer afterwards).
pub struct Keep {
    pub hittables: Vec<RrtHittable>,
    pub children: Vec<i32>,
    pub materials: Vec<RrtMaterial>,
    pub textures: Vec<RrtTexture>,
    pub images: Vec<RrtImage>,
    pub perlins: Vec<RrtPerlin>,
    /// the RGB8 pixels `images[i].rgb` points at (RtwImage::data, rtw_image.rs:57-67), kept alive here
    pub image_pixels: Vec<Arc<Vec<u8>>>,
}

#[derive(Default)]
pub struct SceneFlattener {
    hittables: Vec<RrtHittable>,
    children: Vec<i32>,
    materials: Vec<RrtMaterial>,
    textures: Vec<RrtTexture>,
    images: Vec<RrtImage>,
    perlins: Vec<RrtPerlin>,
    image_pixels: Vec<Arc<Vec<u8>>>,
    seen_hittable: HashMap<*const HittableObject, i32>,
    seen_material: HashMap<*const MaterialObject, i32>,
    seen_texture: HashMap<*const TextureObject, i32>,
}

impl SceneFlattener {
    pub fn new() -> Self {
        Self::default()
    }

    /// One `RrtHittable` per `Arc<HittableObject>`; children are emitted before their parent, a node
    /// reached twice (the glass boundary sphere of the_next_week/mod.rs:558-568) is emitted once.
    pub fn hittable(&mut self, h: &HittableRef) -> i32 {
        if let Some(&i) = self.seen_hittable.get(&Arc::as_ptr(h)) {
            return i;
        }
        let rec = match &**h {
            HittableObject::Sphere(s) => RrtHittable {
                kind: H_SPHERE,
                material: self.material(&s.mat),
                first_child: -1,
                n_children: 0,
                a: s.center.origin().e,    // center0
                b: s.center.direction().e, // center1 - center0 (sphere.rs:28-44)
                c: [s.radius, 0.0, 0.0],
            },
            HittableObject::Quad(q) => RrtHittable {
                kind: H_QUAD,
                material: self.material(&q.mat),
                first_child: -1,
                n_children: 0,
                a: q.q.e,
                b: q.u.e,
                c: q.v.e,
            },
            HittableObject::List(l) => self.container(H_LIST, &l.objects),
            HittableObject::Bvh(b) => self.container(H_BVH, &b.source_objects),
            HittableObject::Translate(t) => {
                let child = self.hittable(&t.object);
                RrtHittable { kind: H_TRANSLATE, material: -1, first_child: child, n_children: 0, a: t.offset.e, ..Default::default() }
            }
            HittableObject::RotateY(r) => {
                let child = self.hittable(&r.object);
                RrtHittable { kind: H_ROTATE_Y, material: -1, first_child: child, n_children: 0, a: [r.angle_degrees, 0.0, 0.0], ..Default::default() }
            }
            HittableObject::ConstantMedium(m) => {
                let boundary = self.hittable(&m.boundary);
                RrtHittable {
                    kind: H_CONSTANT_MEDIUM,
                    material: self.material(&m.phase_function),
                    first_child: boundary,
                    n_children: 0,
                    a: [-1.0 / m.neg_inv_density, 0.0, 0.0], // density (constant_medium.rs:26)
                    ..Default::default()
                }
            }
        };
        self.hittables.push(rec);
        let i = self.hittables.len() as i32 - 1;
        self.seen_hittable.insert(Arc::as_ptr(h), i);
        i
    }

    /// LIST / BVH: the children's indices go into `children[first_child .. first_child + n_children)`.
    /// Children are flattened first, so a nested container's run is complete before the parent's starts.
    fn container(&mut self, kind: i32, objects: &[HittableRef]) -> RrtHittable {
        let ids: Vec<i32> = objects.iter().map(|o| self.hittable(o)).collect();
        let first = self.children.len() as i32;
        self.children.extend_from_slice(&ids);
        RrtHittable { kind, material: -1, first_child: first, n_children: ids.len() as i32, ..Default::default() }
    }

    pub fn material(&mut self, m: &MaterialRef) -> i32 {
        if let Some(&i) = self.seen_material.get(&Arc::as_ptr(m)) {
            return i;
        }
        let rec = match &**m {
            MaterialObject::Lambertian(x) => RrtMaterial { kind: M_LAMBERTIAN, texture: self.texture(&x.tex), ..Default::default() },
            MaterialObject::Metal(x) => RrtMaterial { kind: M_METAL, texture: -1, albedo: x.albedo.e, fuzz: x.fuzz, refraction_index: 0.0 },
            MaterialObject::Dielectric(x) => RrtMaterial { kind: M_DIELECTRIC, texture: -1, refraction_index: x.refraction_index, ..Default::default() },
            MaterialObject::DiffuseLight(x) => RrtMaterial { kind: M_DIFFUSE_LIGHT, texture: self.texture(&x.tex), ..Default::default() },
            MaterialObject::Isotropic(x) => RrtMaterial { kind: M_ISOTROPIC, texture: self.texture(&x.tex), ..Default::default() },
            MaterialObject::EmptyMaterial(_) => RrtMaterial { kind: M_EMPTY, texture: -1, ..Default::default() },
        };
        self.materials.push(rec);
        let i = self.materials.len() as i32 - 1;
        self.seen_material.insert(Arc::as_ptr(m), i);
        i
    }

    pub fn texture(&mut self, t: &TextureRef) -> i32 {
        if let Some(&i) = self.seen_texture.get(&Arc::as_ptr(t)) {
            return i;
        }
        let none = RrtTexture { kind: T_SOLID, even: -1, odd: -1, image: -1, perlin: -1, _pad: 0, color: [0.0; 3], scale: 1.0 };
        let rec = match &**t {
            TextureObject::SolidColor(x) => RrtTexture { color: x.albedo.e, ..none },
            TextureObject::CheckerTexture(x) => {
                let (even, odd) = (self.texture(&x.even), self.texture(&x.odd));
                RrtTexture { kind: T_CHECKER, even, odd, scale: 1.0 / x.inv_scale, ..none } // texture.rs:51
            }
            TextureObject::ImageTexture(x) => RrtTexture { kind: T_IMAGE, image: self.image(&x.image), ..none },
            TextureObject::NoiseTexture(x) => RrtTexture { kind: T_NOISE, perlin: self.perlin(&x.noise), scale: x.scale, ..none },
        };
        self.textures.push(rec);
        let i = self.textures.len() as i32 - 1;
        self.seen_texture.insert(Arc::as_ptr(t), i);
        i
    }

    /// RtwImage after `image::open(..).to_rgb8()` (rtw_image.rs:57-67): row-major RGB8.  A file that was
    /// not found (height 0, rtw_image.rs:34-44) becomes a null image: the library then returns cyan like
    /// texture.rs:91-93.
    pub fn image(&mut self, im: &RtwImage) -> i32 {
        let pixels = Arc::new(im.data.clone());
        let rec = if im.height <= 0 || pixels.is_empty() {
            RrtImage { width: 0, height: 0, rgb: std::ptr::null() }
        } else {
            RrtImage { width: im.width, height: im.height, rgb: pixels.as_ptr() }
        };
        self.image_pixels.push(pixels);
        self.images.push(rec);
        self.images.len() as i32 - 1
    }

    /// Perlin tables as generated by Perlin::new (perlin.rs:11-24): 256 unit vectors + three permutations.
    pub fn perlin(&mut self, p: &Perlin) -> i32 {
        let mut rec = RrtPerlin { randvec: [[0.0; 3]; 256], perm_x: [0; 256], perm_y: [0; 256], perm_z: [0; 256] };
        for i in 0..256 {
            rec.randvec[i] = p.randvec[i].e;
            rec.perm_x[i] = p.perm_x[i] as i32;
            rec.perm_y[i] = p.perm_y[i] as i32;
            rec.perm_z[i] = p.perm_z[i] as i32;
        }
        self.perlins.push(rec);
        self.perlins.len() as i32 - 1
    }

    /// `world`: what the book passes to `Camera::render`; `lights`: book 3's light list (camera.rs:63),
    /// a `HittableList` of proxy geometry with `EmptyMaterial` (mod.rs:128-140), or None for books 1-2.
    pub fn finish(mut self, world: &HittableRef, lights: Option<&HittableRef>) -> (RrtSceneDesc, Keep) {
        let world_id = self.hittable(world);
        let lights_id = lights.map(|l| self.hittable(l)).unwrap_or(-1);
        let keep = Keep {
            hittables: self.hittables,
            children: self.children,
            materials: self.materials,
            textures: self.textures,
            images: self.images,
            perlins: self.perlins,
            image_pixels: self.image_pixels,
        };
        // Vec buffers do not move when `keep` does (a Vec is a pointer to its heap block), so the
        // description stays valid for as long as `keep` lives.
        let desc = RrtSceneDesc {
            hittables: keep.hittables.as_ptr(),
            n_hittables: keep.hittables.len() as i32,
            children: keep.children.as_ptr(),
            n_children: keep.children.len() as i32,
            materials: keep.materials.as_ptr(),
            n_materials: keep.materials.len() as i32,
            textures: keep.textures.as_ptr(),
            n_textures: keep.textures.len() as i32,
            images: keep.images.as_ptr(),
            n_images: keep.images.len() as i32,
            perlins: keep.perlins.as_ptr(),
            n_perlins: keep.perlins.len() as i32,
            world: world_id,
            lights: lights_id,
        };
        (desc, keep)
    }
}
