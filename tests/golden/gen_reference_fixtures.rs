//! gen_reference_fixtures.rs — dump golden vectors from the REAL aligner-core / aligner-helpers.
//!
//! This image has no Rust toolchain, so the oracle of this repository (oracle/*.cpp) is pinned to
//! reference-held vectors only for the global fill (src/tests/test_alignment.rs:14-39).  Everything
//! else — local mode, the stateful del/ext penalty, the arg-max tie rule of ndarray-stats,
//! transform_matrix bit patterns (ndarray::sum order, roots::find_roots_quadratic), the heuristic
//! loop, the PWM aligner — is "parity unpinned".  This file is the one thing that can pin it:
//! run it ONCE on a machine with cargo, commit the JSON it prints, and tests/test_oracle.py
//! (test_reference_generated_fixtures) checks the oracle against every record, bit for bit.
//!
//! How to run (from a checkout of ikramanop/aligner).  Alignment.query / .target are pub(crate)
//! (aligner-core/src/alignment.rs:6-7), so the dump has to live INSIDE the crate as a test module:
//!   cp <this file> aligner-core/src/gen_reference_fixtures.rs
//!   echo '#[cfg(test)] mod gen_reference_fixtures;' >> aligner-core/src/lib.rs
//!   ALNR_FIXTURE_OUT=$PWD/reference_local.json cargo test -p aligner-core --lib gen_reference_fixtures
//!   cp reference_local.json <this repo>/tests/golden/reference_local.json
//! (serde_json, ndarray and aligner-helpers are already dependencies of aligner-core.)
//!
//! Output: one JSON object {"cases": [...], "transform": [...], "heuristic": [...], "pwm": [...]}.
//! f64 values are written as their IEEE-754 bit patterns (u64, decimal) so nothing is lost in text.
//! Every record carries its inputs (sequences, matrix, penalties), so the loader needs nothing else.
//!
//! Reference items exercised:
//!   SimpleGlobalAligner / SimpleLocalAligner::perform_alignment   aligner-core/src/simple/mod.rs:42-144, :168-264
//!   Direction::get_direction(_with_beginning)                     aligner-core/src/enums.rs:18-46
//!   Alignment::get_frequency_matrix / get_alignment               aligner-core/src/alignment.rs:13-42
//!   HeuristicAligner::perform_alignment                           aligner-core/src/heuristic/mod.rs:36-78
//!   PWMAligner::perform_alignment                                 aligner-core/src/pwm/mod.rs:29-126
//!   transform_matrix, get_threshold                               aligner-helpers/src/matrices/mod.rs:8-80
use crate::enums::{BioData, Direction, Protein, DNA};
use crate::heuristic::HeuristicAligner;
use crate::pwm::PWMAligner;
use crate::simple::{SimpleGlobalAligner, SimpleLocalAligner};
use crate::{get_blosum62, AlignerTrait, AlignmentTrait, Heuristics};
use aligner_helpers::matrices::transform_matrix;
use ndarray::{Array1, Array2};
use serde_json::{json, Value};

struct SplitMix(u64);
impl SplitMix {
    fn next(&mut self) -> u64 {
        self.0 = self.0.wrapping_add(0x9E3779B97F4A7C15);
        let mut z = self.0;
        z = (z ^ (z >> 30)).wrapping_mul(0xBF58476D1CE4E5B9);
        z = (z ^ (z >> 27)).wrapping_mul(0x94D049BB133111EB);
        z ^ (z >> 31)
    }
    fn below(&mut self, n: u64) -> u64 { self.next() % n }
    fn unit(&mut self) -> f64 { (self.next() >> 11) as f64 / (1u64 << 53) as f64 }
}

fn bits(v: f64) -> Value { json!(v.to_bits()) }
fn bits2(m: &Array2<f64>) -> Value { json!(m.iter().map(|v| v.to_bits()).collect::<Vec<u64>>()) }
fn dirs(d: &Array2<Direction>) -> Value {
    json!(d.iter().map(|x| match x { Direction::Top => 0, Direction::Left => 1, Direction::Diagonal => 2, Direction::Beginning => 3 }).collect::<Vec<u8>>())
}
fn codes<T: BioData + Into<usize> + Copy>(v: &[T]) -> Vec<usize> { v.iter().map(|s| (*s).into()).collect() }

/// target = query with substitutions (rate 1/4) and a few deletions, alphabet `vol`
fn pair(rng: &mut SplitMix, vol: u64, n: usize) -> (Vec<usize>, Vec<usize>) {
    let q: Vec<usize> = (0..n).map(|_| rng.below(vol) as usize).collect();
    let mut t = Vec::new();
    for &s in &q {
        let u = rng.below(16);
        if u == 0 { continue; }
        t.push(if u < 5 { rng.below(vol) as usize } else { s });
    }
    if t.is_empty() { t.push(q[0]); }
    (q, t)
}

fn simple_case<T>(mode: &str, q: &[usize], t: &[usize], del: f64, ext: f64, m: &Array2<f64>) -> Value
where T: BioData + Into<usize> + From<usize> + Copy + Eq {
    let qs: Vec<T> = q.iter().map(|&c| T::from(c)).collect();
    let ts: Vec<T> = t.iter().map(|&c| T::from(c)).collect();
    macro_rules! dump { ($r:expr) => {{
        let r = $r;
        let a = &r.alignment;
        json!({"mode": mode, "volume": m.shape()[0], "del": bits(del), "ext": bits(ext), "matrix": bits2(m),
               "q": q, "t": t, "H": bits2(&r.alignment_matrix), "D": dirs(&r.direction_matrix),
               "q_aln": codes(&a.query), "t_aln": codes(&a.target), "f": bits(a.f),
               "coords": [a.coords.0 .0, a.coords.0 .1, a.coords.1 .0, a.coords.1 .1],
               "freq": bits2(&a.get_frequency_matrix()), "consensus": codes(&a.get_alignment(m))})
    }}}
    if mode == "global" {
        dump!(SimpleGlobalAligner::<T>::from_seqs(&qs, &ts).unwrap().perform_alignment(del, ext, m, None).unwrap())
    } else {
        dump!(SimpleLocalAligner::<T>::from_seqs(&qs, &ts).unwrap().perform_alignment(del, ext, m, None).unwrap())
    }
}

fn dna_matrix(mat: f64, mis: f64) -> Array2<f64> {
    Array2::from_shape_fn((4, 4), |(i, j)| if i == j { mat } else { mis })
}

#[test]
fn gen_reference_fixtures() {
    let mut rng = SplitMix(0xF1C5_0001);
    let b62 = get_blosum62();
    let mut cases = Vec::new();
    // local mode, del != ext (stateful penalty), ties (small alphabets, +1/-1), both alphabets
    for &(mode, del, ext) in &[("local", 11.0, 2.0), ("local", 11.0, 1.0), ("local", 5.0, 5.0), ("global", 11.0, 2.0), ("global", 8.0, 8.0)] {
        for &n in &[7usize, 33, 150] {
            let (q, t) = pair(&mut rng, 20, n);
            cases.push(simple_case::<Protein>(mode, &q, &t, del, ext, &b62));
        }
    }
    for &(mode, del, ext, mat, mis) in &[("local", 10.0, 2.0, 5.0, -4.0), ("local", 2.0, 1.0, 1.0, -1.0), ("local", 6.0, 6.0, 5.0, -4.0),
                                         ("global", 2.0, 1.0, 1.0, -1.0), ("local", 2.5, 0.75, 5.1, -3.9)] {
        for &n in &[12usize, 60, 150] {
            let (q, t) = pair(&mut rng, 4, n);
            cases.push(simple_case::<DNA>(mode, &q, &t, del, ext, &dna_matrix(mat, mis)));
        }
    }
    // transform_matrix bit patterns (ndarray::sum order, roots 0.0.7, l2_dist)
    let mut transform = Vec::new();
    for k in 0..24 {
        let dim = if k % 3 == 0 { 4 } else { 24 };
        let m = Array2::from_shape_fn((dim, dim), |_| (rng.below(3) as f64) - 1.0 + if k % 2 == 1 { rng.unit() * 0.5 } else { 0.0 });
        let mut fr = Array1::from_shape_fn(dim, |_| 0.05 + rng.unit());
        let s = fr.sum();
        fr.mapv_inplace(|v| v / s);
        let (kd, r2) = (-1.0 + 0.25 * (k % 5) as f64, (dim * dim) as f64 * (1.0 + (k % 4) as f64));
        let out = transform_matrix(&m, kd, r2, &fr);
        transform.push(json!({"matrix": bits2(&m), "dim": dim, "kd": bits(kd), "r2": bits(r2),
                              "freq": fr.iter().map(|v| v.to_bits()).collect::<Vec<u64>>(),
                              "ok": out.is_ok(), "out": out.map(|o| bits2(&o)).unwrap_or(Value::Null)}));
    }
    // heuristic loop: f, matrix of the last pass, strings
    let mut heuristic = Vec::new();
    for k in 0..12 {
        let (q, t) = pair(&mut rng, 20, 60 + 10 * k);
        let m = Array2::from_shape_fn((24, 24), |_| (rng.below(3) as f64) - 1.0);
        let mut cnt = vec![0.0f64; 24];
        for &c in q.iter().chain(t.iter()) { cnt[c] += 1.0; }
        let tot: f64 = cnt.iter().sum();
        let fr = Array1::from_vec(cnt.iter().map(|v| v / tot).collect());
        let qs: Vec<Protein> = q.iter().map(|&c| Protein::from(c)).collect();
        let ts: Vec<Protein> = t.iter().map(|&c| Protein::from(c)).collect();
        let r = HeuristicAligner::<Protein>::from_seqs(&qs, &ts).unwrap()
            .perform_alignment(5.0, 5.0, &m, Some(Heuristics { kd: -1.0, r_squared: 0.0, frequencies: fr.clone() })).unwrap();
        heuristic.push(json!({"q": q, "t": t, "matrix": bits2(&m), "freq": fr.iter().map(|v| v.to_bits()).collect::<Vec<u64>>(),
                              "f": bits(r.alignment.f), "out_matrix": bits2(r.matrix.as_ref().unwrap()),
                              "q_aln": codes(&r.alignment.query), "t_aln": codes(&r.alignment.target),
                              "coords": [r.alignment.coords.0 .0, r.alignment.coords.0 .1, r.alignment.coords.1 .0, r.alignment.coords.1 .1]}));
    }
    // PWM aligner
    let mut pwm = Vec::new();
    for k in 0..8 {
        let l = 20 + 15 * k;
        let prof = Array2::from_shape_fn((4, l), |_| (rng.below(5) as f64) - 2.0 + if k % 2 == 1 { rng.unit() } else { 0.0 });
        let w: Vec<usize> = (0..l + 10).map(|_| rng.below(4) as usize).collect();
        let ws: Vec<DNA> = w.iter().map(|&c| DNA::from(c)).collect();
        let (del, ext) = if k % 2 == 0 { (3.0, 3.0) } else { (4.0, 1.5) };
        let r = PWMAligner::<DNA>::from_seqs(&ws, &[]).unwrap().perform_alignment(del, ext, &prof, None).unwrap();
        pwm.push(json!({"window": w, "profile": bits2(&prof), "length": l, "del": bits(del), "ext": bits(ext),
                        "H": bits2(&r.alignment_matrix), "D": dirs(&r.direction_matrix), "f": bits(r.alignment.f),
                        "numbered": r.alignment.numbered, "query": codes(&r.alignment.query),
                        "coords": [r.alignment.coords.0 .0, r.alignment.coords.0 .1, r.alignment.coords.1 .0, r.alignment.coords.1 .1]}));
    }
    let out = json!({"generator": "tests/golden/gen_reference_fixtures.rs", "cases": cases, "transform": transform,
                     "heuristic": heuristic, "pwm": pwm});
    let path = std::env::var("ALNR_FIXTURE_OUT").unwrap_or_else(|_| "reference_local.json".to_string());
    std::fs::write(&path, out.to_string()).unwrap();
}
