// node.rs — the compute-shader plugin and render-graph node of the reference, served by librt_b200.so.
//
// Replaces, in src/lib.rs of Sur091/GPU-Ray-Tracing:
//   ComputeShaderComputePlugin::build / finish ......... lib.rs:104-136
//   prepare_camera_bind_group / prepare_sphere_buffer /
//   prepare_bind_group / ComputeShaderPipeline ......... lib.rs:144-324   (nothing to prepare: no wgpu pipeline)
//   ComputeShaderState / ComputeShaderNode ............. lib.rs:326-422
// and adds `present_frame`, which puts the finished frame where the sprite samples it (lib.rs:79-102): without
// GL / Vulkan interop the texels travel through host memory, overlapped with the next frame's kernel.
// Everything else of the app (window, sprite, camera controller, scene generation) is untouched; `SIZE`,
// `ComputeShaderImages`, `camera::SceneCamera` and `scene::sphere::SphereCollection` are the reference's own items.
//
// Not compiled in the repository's build image (no Rust toolchain there).
use crate::renderer::Renderer;
use crate::scene::sphere::SphereCollection;
use crate::{camera, ComputeShaderImages, SIZE};
use bevy::{
    prelude::*,
    render::{
        extract_resource::ExtractResourcePlugin,
        render_graph::{self, RenderGraph, RenderLabel},
        renderer::RenderContext,
        RenderApp,
    },
};
use std::sync::{Arc, Mutex};

pub struct ComputeShaderComputePlugin;

#[derive(Debug, Hash, PartialEq, Eq, Clone, RenderLabel)]
struct ComputeShaderLabel;

/// The device-side state of the pass; lives in the render world, where the bind groups lived.
#[derive(Resource)]
struct B200Renderer(Renderer);

/// Frame hand-off render world -> main world: the texels of the last finished frame (rgba32float, as the
/// storage texture held them).  `Arc<Mutex<..>>` because the two worlds may run on different threads.
#[derive(Resource, Clone, Default)]
pub struct FinishedFrame(pub Arc<Mutex<Vec<f32>>>);

impl Plugin for ComputeShaderComputePlugin {
    fn build(&self, app: &mut App) {
        let frame = FinishedFrame::default();
        // the same three resources cross into the render world as before (lib.rs:113-116, sphere.rs:162)
        app.add_plugins((
            ExtractResourcePlugin::<ComputeShaderImages>::default(),
            ExtractResourcePlugin::<camera::SceneCamera>::default(),
        ))
        .insert_resource(frame.clone())
        .add_systems(Update, present_frame);

        let render_app = app.sub_app_mut(RenderApp);
        render_app.insert_resource(frame);
        let mut render_graph = render_app.world_mut().resource_mut::<RenderGraph>();
        render_graph.add_node(ComputeShaderLabel, ComputeShaderNode::default());
        // the pass still runs before the sprite is drawn (lib.rs:127-129)
        render_graph.add_node_edge(ComputeShaderLabel, bevy::render::graph::CameraDriverLabel);
    }

    fn finish(&self, app: &mut App) {
        // where ComputeShaderPipeline::from_world compiled the shader (lib.rs:132-135, 240-324): create the handle;
        // the kernels are in the library, so there is no Loading phase to wait for
        let renderer = Renderer::new(0, SIZE.0, SIZE.1).unwrap_or_else(|e| panic!("creating the B200 renderer: {e}"));
        app.sub_app_mut(RenderApp).insert_resource(B200Renderer(renderer));
    }
}

enum ComputeShaderState {
    Loading,
    Init,
    Update(usize),
}

struct ComputeShaderNode {
    state: ComputeShaderState,
}

impl Default for ComputeShaderNode {
    fn default() -> Self {
        Self { state: ComputeShaderState::Loading }
    }
}

impl render_graph::Node for ComputeShaderNode {
    fn update(&mut self, _world: &mut World) {
        // same machine as lib.rs:345-377; the ping-pong index is carried for parity only (the accumulation image is
        // updated in place: every invocation touches its own texel, compute_shader.wgsl:339, 363)
        self.state = match self.state {
            ComputeShaderState::Loading => ComputeShaderState::Init,
            ComputeShaderState::Init => ComputeShaderState::Update(1),
            ComputeShaderState::Update(0) => ComputeShaderState::Update(1),
            ComputeShaderState::Update(_) => ComputeShaderState::Update(0),
        };
    }

    fn run(
        &self,
        _graph: &mut render_graph::RenderGraphContext,
        _render_context: &mut RenderContext,
        world: &World,
    ) -> Result<(), render_graph::NodeRunError> {
        let renderer = &world.resource::<B200Renderer>().0;
        let outcome = match self.state {
            ComputeShaderState::Loading => Ok(()),
            ComputeShaderState::Init => {
                // prepare_sphere_buffer's upload (lib.rs:177-207) + the `init` dispatch (lib.rs:398-407)
                let spheres = world.resource::<SphereCollection>();
                renderer.set_scene(&spheres.spheres, spheres.count).and_then(|_| renderer.init())
            }
            ComputeShaderState::Update(_) => {
                // prepare_camera_bind_group's 176 bytes (lib.rs:151-175) + the `update` dispatch (lib.rs:408-417).
                // The reference re-uploads the sphere buffer every frame as well; its scene is static
                // (sphere.rs:167-170), and an upload of identical bytes would keep the tree the library built.
                let scene_camera = world.resource::<camera::SceneCamera>();
                renderer.update(scene_camera).and_then(|_| {
                    let frame = world.resource::<FinishedFrame>();
                    let mut texels = frame.0.lock().unwrap();
                    texels.resize((SIZE.0 * SIZE.1 * 4) as usize, 0.0);
                    renderer.read_rgba32f(&mut texels)
                })
            }
        };
        // the reference panics / unwraps at the corresponding places (lib.rs:216-217, 356-358, 399-401)
        outcome.unwrap_or_else(|e| panic!("{e}"));
        Ok(())
    }
}

/// Main world: copy the finished frame into the image the sprite shows.  The format stays Rgba32Float
/// (lib.rs:63-75), so the bytes are the texels as the shader's textureStore wrote them (linear, alpha = count).
fn present_frame(frame: Res<FinishedFrame>, handles: Res<ComputeShaderImages>, mut images: ResMut<Assets<Image>>) {
    let texels = frame.0.lock().unwrap();
    if texels.is_empty() {
        return;
    }
    // switch_textures (lib.rs:96-102) alternates the sprite between a and b every frame: keep both current
    for handle in [&handles.texture_a, &handles.texture_b] {
        if let Some(image) = images.get_mut(handle) {
            image.data.clear();
            image.data.extend_from_slice(bytemuck::cast_slice(&texels));
        }
    }
}
