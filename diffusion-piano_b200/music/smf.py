"""Standard MIDI File reader/writer -> NoteSequence, with no third-party dependency.

Replaces `note_seq.midi_io.midi_file_to_note_sequence` as called by the reference at
`robopianist/music/midi_file.py:174-178` and `data_processing/add_fingering_to_midi.py:64`.
note_seq delegates the parse to pretty_midi; this module restates pretty_midi's published
loading rules so that note order, note pairing and tick->seconds arithmetic come out the
same (they matter: `MidiFile.trim_silence` uses *list order* notes[0]/notes[-1],
midi_file.py:231-237, and frame indices are `int(t * fps)`, piano_roll.py:97):

* tempo map is read from track 0 only; tick_scale = 60 / ((6e7 / tempo_us) * resolution);
  time(tick) = time(segment start) + tick_scale * (tick - segment start);
* a note-off closes *every* open note-on of that (channel, pitch) in the track whose start
  tick differs from the note-off tick; a note is appended to its instrument when it closes;
* instruments are keyed (program, channel, track) in order of first *closed note*; control
  changes on a (channel, track) that never produces a note are dropped;
* NoteSequence.total_time = max note end; notes listed instrument by instrument.
"""

from __future__ import annotations

import struct
from collections import OrderedDict, defaultdict
from pathlib import Path
from typing import List, Tuple, Union

from .sequence import NoteSequence


class MidiParseError(RuntimeError):
    pass


def _read_varlen(data: bytes, i: int) -> Tuple[int, int]:
    value = 0
    while True:
        if i >= len(data):
            raise MidiParseError("truncated variable-length quantity")
        b = data[i]
        i += 1
        value = (value << 7) | (b & 0x7F)
        if not b & 0x80:
            return value, i


def _parse_track(data: bytes) -> List[tuple]:
    """Returns a list of (abs_tick, kind, ...) events."""
    events = []
    i = 0
    tick = 0
    status = None
    n = len(data)
    while i < n:
        delta, i = _read_varlen(data, i)
        tick += delta
        if i >= n:
            break
        b = data[i]
        if b == 0xFF:  # meta
            mtype = data[i + 1]
            length, j = _read_varlen(data, i + 2)
            payload = data[j : j + length]
            i = j + length
            if mtype == 0x51 and length == 3:
                events.append((tick, "set_tempo", int.from_bytes(payload, "big")))
            elif mtype == 0x58 and length >= 2:
                events.append((tick, "time_signature", payload[0], 1 << payload[1]))
            elif mtype == 0x2F:
                events.append((tick, "end_of_track"))
            else:
                events.append((tick, "meta", mtype))
            continue
        if b in (0xF0, 0xF7):  # sysex
            length, j = _read_varlen(data, i + 1)
            i = j + length
            events.append((tick, "sysex"))
            continue
        if b & 0x80:
            status = b
            i += 1
        elif status is None:
            raise MidiParseError("data byte without running status")
        hi = status & 0xF0
        ch = status & 0x0F
        if hi in (0x80, 0x90, 0xA0, 0xB0, 0xE0):
            d1, d2 = data[i], data[i + 1]
            i += 2
            if hi == 0x90:
                events.append((tick, "note_on", ch, d1, d2))
            elif hi == 0x80:
                events.append((tick, "note_off", ch, d1, d2))
            elif hi == 0xB0:
                events.append((tick, "control_change", ch, d1, d2))
            else:
                events.append((tick, "other"))
        elif hi in (0xC0, 0xD0):
            d1 = data[i]
            i += 1
            if hi == 0xC0:
                events.append((tick, "program_change", ch, d1))
            else:
                events.append((tick, "other"))
        else:  # system common/realtime inside a track: skip conservatively
            if status == 0xF2:
                i += 2
            elif status == 0xF3:
                i += 1
            events.append((tick, "other"))
    return events


def parse_smf_events(data: bytes):
    """bytes -> (resolution, tracks) where tracks[i] is a list of absolute-tick events."""
    if data[:4] != b"MThd":
        raise MidiParseError("not a Standard MIDI File (missing MThd)")
    hlen = struct.unpack(">I", data[4:8])[0]
    fmt, ntrk, division = struct.unpack(">HHH", data[8:14])
    if division & 0x8000:
        raise MidiParseError("SMPTE time division is not supported")
    tracks = []
    i = 8 + hlen
    while i + 8 <= len(data) and len(tracks) < ntrk:
        tag = data[i : i + 4]
        length = struct.unpack(">I", data[i + 4 : i + 8])[0]
        body = data[i + 8 : i + 8 + length]
        i += 8 + length
        if tag == b"MTrk":
            tracks.append(_parse_track(body))
    if not tracks:
        raise MidiParseError("no tracks")
    return division, tracks


def parse_smf(data: bytes) -> NoteSequence:
    resolution, tracks = parse_smf_events(data)
    return events_to_sequence(resolution, tracks)


def events_to_sequence(resolution: int, tracks) -> NoteSequence:
    """pretty_midi's loading rules + note_seq's assembly order (see module docstring)."""
    # --- tempo map (track 0 only), pretty_midi._load_tempo_changes ---
    tick_scales = [(0, 60.0 / (120.0 * resolution))]
    for ev in tracks[0]:
        if ev[1] != "set_tempo":
            continue
        tick, tempo_us = ev[0], ev[2]
        if tick == 0:
            bpm = 6e7 / tempo_us
            tick_scales = [(0, 60.0 / (bpm * resolution))]
        else:
            last = tick_scales[-1][1]
            scale = 60.0 / ((6e7 / tempo_us) * resolution)
            if scale != last:
                tick_scales.append((tick, scale))

    seg_start_time = [0.0]
    for (s0, sc0), (s1, _) in zip(tick_scales[:-1], tick_scales[1:]):
        seg_start_time.append(seg_start_time[-1] + sc0 * (s1 - s0))

    def tick_to_time(tick: int) -> float:
        k = 0
        for idx in range(len(tick_scales)):
            if tick_scales[idx][0] <= tick:
                k = idx
        # a tick exactly on a boundary belongs to the *earlier* segment's formula in
        # pretty_midi (it writes [start:end+1] and then the next segment overwrites from
        # its own start with last_end_time + 0) - identical value either way.
        s, sc = tick_scales[k]
        return seg_start_time[k] + sc * (tick - s)

    seq = NoteSequence()
    seq.ticks_per_quarter = resolution
    for k, (tick, scale) in enumerate(tick_scales):
        seq.tempos.add(time=tick_to_time(tick), qpm=60.0 / (scale * resolution))
    for ev in tracks[0]:
        if ev[1] == "time_signature":
            seq.time_signatures.add(
                time=tick_to_time(ev[0]), numerator=ev[2], denominator=ev[3]
            )

    # --- instruments, pretty_midi._load_instruments ---
    instrument_map: "OrderedDict[tuple, dict]" = OrderedDict()
    stragglers = {}

    def get_instrument(program, channel, track, create_new):
        key = (program, channel, track)
        if key in instrument_map:
            return instrument_map[key]
        if not create_new and (channel, track) in stragglers:
            return stragglers[(channel, track)]
        inst = {"program": program, "is_drum": channel == 9, "notes": [], "ccs": []}
        if create_new:
            if (channel, track) in stragglers:
                inst["ccs"] = stragglers[(channel, track)]["ccs"]
            instrument_map[key] = inst
        else:
            stragglers[(channel, track)] = inst
        return inst

    for track_idx, track in enumerate(tracks):
        last_note_on = defaultdict(list)
        current_program = [0] * 16
        for ev in track:
            kind = ev[1]
            if kind == "program_change":
                current_program[ev[2]] = ev[3]
            elif kind == "note_on" and ev[4] > 0:
                last_note_on[(ev[2], ev[3])].append((ev[0], ev[4]))
            elif kind == "note_off" or (kind == "note_on" and ev[4] == 0):
                key = (ev[2], ev[3])
                if key in last_note_on:
                    end_tick = ev[0]
                    open_notes = last_note_on[key]
                    to_close = [(s, v) for s, v in open_notes if s != end_tick]
                    to_keep = [(s, v) for s, v in open_notes if s == end_tick]
                    for s, v in to_close:
                        inst = get_instrument(current_program[ev[2]], ev[2], track_idx, True)
                        inst["notes"].append(
                            (ev[3], v, tick_to_time(s), tick_to_time(end_tick))
                        )
                    if to_close and to_keep:
                        last_note_on[key] = to_keep
                    else:
                        del last_note_on[key]
            elif kind == "control_change":
                inst = get_instrument(current_program[ev[2]], ev[2], track_idx, False)
                inst["ccs"].append((ev[3], ev[4], tick_to_time(ev[0])))

    # --- note_seq.midi_io.midi_to_note_sequence assembly order ---
    for num, inst in enumerate(instrument_map.values()):
        for pitch, vel, start, end in inst["notes"]:
            if not seq.total_time or end > seq.total_time:
                seq.total_time = end
            seq.notes.add(
                pitch=pitch, velocity=vel, start_time=start, end_time=end,
                instrument=num, program=inst["program"], is_drum=inst["is_drum"],
            )
    for num, inst in enumerate(instrument_map.values()):
        for number, value, time in inst["ccs"]:
            seq.control_changes.add(
                time=time, control_number=number, control_value=value,
                instrument=num, program=inst["program"], is_drum=inst["is_drum"],
            )
    return seq


def read_midi(path: Union[str, Path]) -> NoteSequence:
    return parse_smf(Path(path).read_bytes())


# ----------------------------------------------------------------------------------
# writer (used by tests and by MidiFile.save); one tempo, one track of notes + CC64.
# ----------------------------------------------------------------------------------
def _varlen(v: int) -> bytes:
    out = [v & 0x7F]
    v >>= 7
    while v:
        out.append((v & 0x7F) | 0x80)
        v >>= 7
    return bytes(reversed(out))


def write_smf(seq: NoteSequence, ticks_per_quarter: int = 480) -> bytes:
    qpm = seq.tempos[0].qpm if seq.tempos else 120.0
    tempo_us = int(round(6e7 / qpm))
    sec_per_tick = 60.0 / ((6e7 / tempo_us) * ticks_per_quarter)

    def to_tick(t: float) -> int:
        return int(round(t / sec_per_tick))

    meta = _varlen(0) + b"\xff\x51\x03" + tempo_us.to_bytes(3, "big") + _varlen(0) + b"\xff\x2f\x00"
    evs = []
    for n in seq.notes:
        evs.append((to_tick(n.start_time), 1, bytes([0x90, n.pitch, max(1, n.velocity)])))
        evs.append((to_tick(n.end_time), 0, bytes([0x80, n.pitch, 0])))
    for c in seq.control_changes:
        evs.append((to_tick(c.time), 2, bytes([0xB0, c.control_number, c.control_value])))
    evs.sort(key=lambda e: (e[0], e[1]))
    body = b""
    last = 0
    for tick, _, raw in evs:
        body += _varlen(tick - last) + raw
        last = tick
    body += _varlen(0) + b"\xff\x2f\x00"
    out = b"MThd" + struct.pack(">IHHH", 6, 1, 2, ticks_per_quarter)
    out += b"MTrk" + struct.pack(">I", len(meta)) + meta
    out += b"MTrk" + struct.pack(">I", len(body)) + body
    return out
