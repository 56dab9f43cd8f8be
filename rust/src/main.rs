// main.rs — same as the reference's (src/main.rs:1-3): `cargo run --release` renders output.ppm.
fn main() {
    nerf_rs::render_cli_image();
}
