// UNBUILT here: this image has no rustc/cargo.  Kept in sync with include/rrt.h by
// tests/test_abi.py::test_rust_binding_declares_every_entry_point.  See INTEGRATION.md.
// Visitor over the reference's HittableObject tree -> rrt_scene_desc arrays (sketch; the Rust structs
// need `source_objects` on BvhNode and `angle_degrees` on RotateY).
impl SceneFlattener {
    fn hittable(&mut self, h: &HittableRef) -> i32 {
        if let Some(&i) = self.seen.get(&Arc::as_ptr(h)) { return i; }
        let rec = match &**h {
            HittableObject::Sphere(s)  => RrtHittable { kind: 0, material: self.material(&s.mat), first_child: -1, n_children: 0,
                                                        a: s.center.origin().e, b: s.center.direction().e, c: [s.radius, 0.0, 0.0] },
            HittableObject::Quad(q)    => RrtHittable { kind: 1, material: self.material(&q.mat), first_child: -1, n_children: 0,
                                                        a: q.q.e, b: q.u.e, c: q.v.e },
            HittableObject::List(l)    => self.container(2, &l.objects),
            HittableObject::Bvh(b)     => self.container(3, &b.source_objects),   // keep the list BvhNode::new was given
            HittableObject::Translate(t) => { let c = self.hittable(&t.object);
                                              RrtHittable { kind: 4, material: -1, first_child: c, n_children: 0, a: t.offset.e, ..Default::default() } }
            HittableObject::RotateY(r) => { let c = self.hittable(&r.object);
                                            RrtHittable { kind: 5, material: -1, first_child: c, n_children: 0, a: [r.angle_degrees, 0.0, 0.0], ..Default::default() } }
            HittableObject::ConstantMedium(m) => { let c = self.hittable(&m.boundary);
                                            RrtHittable { kind: 6, material: self.material(&m.phase_function), first_child: c, n_children: 0,
                                                          a: [-1.0 / m.neg_inv_density, 0.0, 0.0], ..Default::default() } }
        };
        self.hittables.push(rec); let i = self.hittables.len() as i32 - 1; self.seen.insert(Arc::as_ptr(h), i); i
    }
}
