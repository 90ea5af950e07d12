//! Replacement body for `Scene::draw_scene` (src/scene/engine.rs:186-255): flattens the RayTracer the
//! reference already built (triangles in `octree.triangles` order == `f`-line order, src/collision/octree.rs:45-47)
//! and renders the whole frame on the GPU into `self.canvas.buffer`.  NOT COMPILED HERE (no Rust toolchain).
//!
//! Paste into `impl Scene` in src/scene/engine.rs (the canvas `buffer` field must be visible to it).
use crate::rrt_ffi::*;
use crate::scene::entities::Light;
use std::collections::HashMap;
use std::sync::Arc;

pub fn draw_scene(&mut self, rt: RayTracer) {
    let tris = &rt.scene_data.octree.triangles;
    let n = tris.len();
    let (mut v, mut vt, mut vn) = (Vec::with_capacity(9 * n), Vec::with_capacity(6 * n), Vec::with_capacity(9 * n));
    let mut material_id: Vec<i32> = Vec::with_capacity(n);
    // materials / textures are shared through Arc: index them by pointer identity
    let mut tex_index: HashMap<*const crate::scene::entities::Texture, i32> = HashMap::new();
    let mut tex_bytes: Vec<Vec<u8>> = vec![];
    let mut tex_dims: Vec<(i32, i32)> = vec![];
    let mut mat_index: HashMap<*const crate::scene::material::Material, i32> = HashMap::new();
    let mut mats: Vec<rrt_material> = vec![];
    let mut intern_tex = |t: &Arc<crate::scene::entities::Texture>| -> i32 {
        *tex_index.entry(Arc::as_ptr(t)).or_insert_with(|| {
            tex_bytes.push(t.colours.iter().flat_map(|c| [c.r, c.g, c.b]).collect());
            tex_dims.push((t.width as i32, t.height as i32));
            (tex_bytes.len() - 1) as i32
        })
    };
    for t in tris {
        for p in [&t.v1, &t.v2, &t.v3] { v.extend_from_slice(&[p.x, p.y, p.z]); }
        for p in [&t.v1_tex_coords, &t.v2_tex_coords, &t.v3_tex_coords] { vt.extend_from_slice(&[p.x, p.y]); }
        for p in [&t.v1_normal_coords, &t.v2_normal_coords, &t.v3_normal_coords] { vn.extend_from_slice(&[p.x, p.y, p.z]); }
        let m = &t.material;
        let key = Arc::as_ptr(m);
        if !mat_index.contains_key(&key) {
            let tex_id = intern_tex(&m.texture);
            let bump_id = m.bump_map.as_ref().map(|b| intern_tex(b)).unwrap_or(-1);
            let a = m.ambient_color_coefficient; let d = m.diffuse_color_coefficient; let s = m.specular_color_coefficient;
            mats.push(rrt_material { ka: [a.x, a.y, a.z], kd: [d.x, d.y, d.z], ks: [s.x, s.y, s.z],
                                     ns: m.specular_weight, kr: m.reflectivity, tex_id, bump_id });
            mat_index.insert(key, (mats.len() - 1) as i32);
        }
        material_id.push(mat_index[&key]);
    }
    let textures: Vec<rrt_texture> = tex_bytes.iter().zip(&tex_dims)
        .map(|(b, (w, h))| rrt_texture { width: *w, height: *h, rgb: b.as_ptr() }).collect();
    let root = &rt.scene_data.octree.aabbs[0]; // the root box (src/file_management/utils.rs:145)
    let desc = rrt_scene_desc {
        n_tris: n as u64, v: v.as_ptr(), vt: vt.as_ptr(), vn: vn.as_ptr(), material_id: material_id.as_ptr(),
        n_materials: mats.len() as u32, materials: mats.as_ptr(),
        n_textures: textures.len() as u32, textures: textures.as_ptr(),
        root_min: [root.min_coords.x, root.min_coords.y, root.min_coords.z],
        root_max: [root.max_coords.x, root.max_coords.y, root.max_coords.z],
    };
    unsafe {
        let mut p: rrt_render_params = std::mem::zeroed();
        rrt_render_params_default(&mut p);
        p.width = self.canvas.width as i32;
        p.height = self.canvas.height as i32;
        p.origin = [rt.origin.x, rt.origin.y, rt.origin.z];
        p.viewport = [self.viewport.width, self.viewport.height, self.viewport.distance];
        p.n_lights = rt.lights.len() as i32;
        for (i, l) in rt.lights.iter().enumerate() {
            p.lights[i] = match l {
                Light::Ambient { intensity } => rrt_light { kind: 0, _pad: 0, intensity: *intensity, vec: [0.0; 3] },
                Light::Point { intensity, position } => rrt_light { kind: 1, _pad: 0, intensity: *intensity, vec: [position.x, position.y, position.z] },
                Light::Directional { intensity, direction } => rrt_light { kind: 2, _pad: 0, intensity: *intensity, vec: [direction.x, direction.y, direction.z] },
            };
        }
        let mut scene: *mut rrt_scene = std::ptr::null_mut();
        if rrt_scene_create(&desc, 0, &mut scene) != 0 { panic!("rrt_scene_create: {}", last_error()); }
        // canvas.buffer is a pageable Vec<u32>: rrt_render renders into device memory and copies.  Page-lock it once
        // (cudaHostRegister on buffer.as_ptr(), outside this function) and the kernel stores the pixels into it directly.
        let rc = rrt_render(scene, &p, self.canvas.buffer.as_mut_ptr(), std::ptr::null_mut(), std::ptr::null_mut());
        rrt_scene_destroy(scene);
        if rc != 0 { panic!("rrt_render: {}", last_error()); }
    }
    self.canvas.update(); // one blit instead of one per 50-row chunk (engine.rs:253)
}
