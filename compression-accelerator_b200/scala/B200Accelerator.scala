// B200Accelerator.scala -- JNA binding of libsnappy_b200's RoCC device object (include/b200_rocc.h)
// and a PeekPokeTester-shaped facade, so that the reference's integration test
// (src/test/scala/example/Test.scala:17-79, CompressionAcceleratorTester) can drive the B200
// library with the same poke / step / peek calls it uses on the Chisel ScratchpadTestModule
// (src/main/scala/example/ScratchpadTest.scala:23-51).
//
// NOT COMPILED OR RUN in this repository's environment: neither the build container nor the GPU
// box has a JVM / scalac / sbt (SURVEY.md section 8c).  The C side it binds IS tested, through
// the same protocol, by tests/test_gpu_parity.py::test_rocc_device_protocol.
//
// sbt: libraryDependencies += "net.java.dev.jna" % "jna" % "5.13.0"
// run: -Djna.library.path=<repo>/compression-accelerator_b200
package example

import com.sun.jna.{Library, Native, Pointer}

trait SnappyB200Lib extends Library {
  // include/b200_rocc.h
  def b200_rocc_new(memBytes: Long, cudaDevice: Int): Pointer
  def b200_rocc_free(dev: Pointer): Unit
  def b200_rocc_load_mem(dev: Pointer, bytes: Array[Byte], n: Long, addr: Long): Long
  def b200_rocc_load_mem_file(dev: Pointer, path: String, addr: Long): Long
  def b200_rocc_mem(dev: Pointer): Pointer
  def b200_rocc_mem_bytes(dev: Pointer): Long
  def b200_rocc_poke_cmd(dev: Pointer, funct: Int, rs1: Long, rs2: Long, valid: Int): Unit
  def b200_rocc_step(dev: Pointer, n: Long): Unit
  def b200_rocc_peek_cmd_ready(dev: Pointer): Int
  def b200_rocc_peek_busy(dev: Pointer): Int
  def b200_rocc_peek_resp_valid(dev: Pointer): Int
  def b200_rocc_peek_resp_data(dev: Pointer): Long
  def b200_rocc_peek_busy_cycles(dev: Pointer): Long
  def b200_rocc_peek_interrupt(dev: Pointer): Int
  // include/snappy_b200.h (driver contract of tests/compression.c:20-39)
  def snappy_b200_rocc_compress(source: Array[Byte], uncompressedLength: Int, dest: Array[Byte]): Long
  def snappy_b200_max_compressed_length(n: Long): Long
  def snappy_b200_last_error(): String
}

object SnappyB200Lib {
  lazy val instance: SnappyB200Lib = Native.load("snappy_b200", classOf[SnappyB200Lib])
}

/** Signals of ScratchpadTestModule.io that the tester touches, as plain handles. */
sealed trait B200Signal
object B200Signal {
  case object CmdFunct extends B200Signal     // c.io.cmd.bits.inst.funct
  case object CmdRs1 extends B200Signal       // c.io.cmd.bits.rs1
  case object CmdRs2 extends B200Signal       // c.io.cmd.bits.rs2
  case object CmdValid extends B200Signal     // c.io.cmd.valid
  case object CmdReady extends B200Signal     // c.io.cmd.ready
  case object RespValid extends B200Signal    // c.io.resp.valid
  case object RespData extends B200Signal     // c.io.resp.bits.data
  case object Busy extends B200Signal         // c.io.busy
  case object Interrupt extends B200Signal    // c.io.interrupt
  case object BusyCycles extends B200Signal   // c.io.busyCycles
}

/** Stands where `LazyModule(new ScratchpadTest(OpcodeSet.custom3, filename)).module` stands in
  * Test.scala:68: memory preloaded from `filename` at address 0. */
class B200ScratchpadTestModule(filename: String, cudaDevice: Int = 0) {
  private val lib = SnappyB200Lib.instance
  val dev: Pointer = lib.b200_rocc_new(0L, cudaDevice)
  require(dev != null, "b200_rocc_new failed: " + lib.snappy_b200_last_error())
  if (filename.nonEmpty) require(lib.b200_rocc_load_mem_file(dev, filename, 0L) >= 0, "cannot load " + filename)
  private[example] var funct = 0
  private[example] var rs1 = 0L
  private[example] var rs2 = 0L
  private[example] var valid = 0
  def close(): Unit = lib.b200_rocc_free(dev)
  def memory(addr: Long, n: Int): Array[Byte] = lib.b200_rocc_mem(dev).getByteArray(addr, n)
}

/** poke / peek / step with the meaning chisel3.iotesters.PeekPokeTester gives them. */
class B200PeekPokeTester(val c: B200ScratchpadTestModule) {
  private val lib = SnappyB200Lib.instance
  import B200Signal._
  def poke(sig: B200Signal, value: BigInt): Unit = {
    sig match {
      case CmdFunct => c.funct = value.toInt
      case CmdRs1   => c.rs1 = value.toLong
      case CmdRs2   => c.rs2 = value.toLong
      case CmdValid => c.valid = if (value != 0) 1 else 0
      case other    => throw new IllegalArgumentException("not an input: " + other)
    }
    lib.b200_rocc_poke_cmd(c.dev, c.funct, c.rs1, c.rs2, c.valid)
  }
  def poke(sig: B200Signal, value: Boolean): Unit = poke(sig, if (value) BigInt(1) else BigInt(0))
  def peek(sig: B200Signal): BigInt = sig match {
    case CmdReady   => BigInt(lib.b200_rocc_peek_cmd_ready(c.dev))
    case RespValid  => BigInt(lib.b200_rocc_peek_resp_valid(c.dev))
    case RespData   => BigInt(lib.b200_rocc_peek_resp_data(c.dev))
    case Busy       => BigInt(lib.b200_rocc_peek_busy(c.dev))
    case Interrupt  => BigInt(lib.b200_rocc_peek_interrupt(c.dev))
    case BusyCycles => BigInt(lib.b200_rocc_peek_busy_cycles(c.dev))
    case other      => throw new IllegalArgumentException("not an output: " + other)
  }
  def step(n: Int): Unit = lib.b200_rocc_step(c.dev, n.toLong)
}

/** CompressionAcceleratorTester (Test.scala:17-54) with c.io.* replaced by the handles above;
  * the statement order, the constants and the printed line are the reference's. */
class B200CompressionAcceleratorTester(c: B200ScratchpadTestModule, filename: String) extends B200PeekPokeTester(c) {
  import B200Signal._
  var timeout = 0
  val length: Int = filename.split("-").last.split(".txt")(0).toInt
  val dataset: String = filename.split("/").last.split(".txt")(0).split("-")(0)
  println("== type: " + dataset + "; length: " + length + ";")
  poke(CmdFunct, 2)                  // doSetLength
  poke(CmdRs1, length)
  poke(CmdValid, true)
  step(1)
  poke(CmdFunct, 0)                  // doCompress
  poke(CmdRs1, 0)                    // src = 0
  poke(CmdRs2, (Math.ceil(length / 8.0) * 8.0).toInt + 8)
  step(1)
  poke(CmdValid, false)
  while (peek(RespValid) == 0 && timeout < 50 * length) {
    step(1)
    timeout += 1
  }
  step(20)
  println("== cycles: " + peek(BusyCycles) + "; compressed length: " + peek(RespData))
}

object B200CompressionAcceleratorSpec extends App {
  val dir = new java.io.File("benchmark/benchmark-data/")
  val files = dir.listFiles().filter(!_.getName.contains("_")).filter(!_.getName.contains("gitkeep"))
  for (filename <- files.map(_.toString)) {
    val m = new B200ScratchpadTestModule(filename)
    try new B200CompressionAcceleratorTester(m, filename) finally m.close()
  }
}
